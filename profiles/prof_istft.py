"""Driver for ncu: ISTFT on the cfg2 batch shape."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import speechbrain_b200 as sb
wav = torch.randn(32, 256000, device="cuda") * 0.1
spec = sb.STFT(sample_rate=16000)(wav)
ist = sb.ISTFT(sample_rate=16000)
for _ in range(3):
    out = ist(spec)
torch.cuda.synchronize()
print("ok")
