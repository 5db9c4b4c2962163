"""Times Fbank(n_fft=512, n_mels=80) on the cfg2 batch shape (generic path: power-of-two FFT kernel
+ mel kernel + clamp).  python profiles/nfft512_bench.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import speechbrain_b200 as sb

torch.manual_seed(0)
wav = torch.randn(32, 256000, device="cuda") * 0.1
for n_fft in (512, 400):
    fb = sb.Fbank(n_mels=80, n_fft=n_fft)
    for _ in range(3):
        out = fb(wav)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        out = fb(wav)
    e1.record()
    torch.cuda.synchronize()
    print("Fbank(n_fft=%d, n_mels=80) 32 x 16 s: %.1f us per call" % (n_fft, e0.elapsed_time(e1) * 100))
