"""Driver for ncu: Fbank(n_fft=512, n_mels=80) on the cfg2 batch shape (general path)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import speechbrain_b200 as sb
wav = torch.randn(32, 256000, device="cuda") * 0.1
fb = sb.Fbank(n_mels=80, n_fft=512)
for _ in range(3):
    out = fb(wav)
torch.cuda.synchronize()
print("ok")
