"""Driver for ncu: the 512-point fused kernel and the tensor-core projection on the cfg2 batch."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import speechbrain_b200 as sb

torch.manual_seed(0)
wav = torch.randn(32, 256000, device="cuda") * 0.1
fb512 = sb.Fbank(n_mels=80, n_fft=512)
gauss = sb.Fbank(n_mels=80, n_fft=512, filter_shape="gaussian")     # STFT + mel_mma_kernel
for _ in range(3):
    a = fb512(wav)
    b = gauss(wav)
torch.cuda.synchronize()
print("ok", float(a.mean()), float(b.mean()))
