"""ncu driver: SpecAugment kernel on the cfg2 feature tensor (warp + masks, then masks only)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from speechbrain_b200 import ops
B, T, F = 32, 1601, 80
torch.manual_seed(0)
x = torch.randn(B, T, F, device="cuda")
fm = (torch.randint(0, 60, (B, 2), device="cuda"), torch.randint(0, 20, (B, 2), device="cuda"))
tm = (torch.randint(0, 1500, (B, 2), device="cuda"), torch.randint(0, 100, (B, 2), device="cuda"))
for _ in range(2):
    y = ops.specaugment_forward(x, (800, 803), "bicubic", fm, tm, 0.0)
    ops.specaugment_forward(x.clone(), None, "bicubic", fm, tm, 0.0)
torch.cuda.synchronize()
print("ok", float(y.abs().mean()))
