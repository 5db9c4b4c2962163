"""Time-domain augmentation on the cfg2 batch (32 x 16 s): this repo's modules (host draws + one
kernel each), CUDA events, eager.   python profiles/tdbench.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import speechbrain_b200 as sb

torch.manual_seed(0)
B, L = 32, 256000
wav = torch.randn(B, L, device="cuda") * 0.1
lens = torch.linspace(0.6, 1.0, B, device="cuda")
BYTES = 2 * 4 * B * L


def timed(fn, n=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3


mods = [("DropChunk", sb.DropChunk(), True), ("DropFreq", sb.DropFreq(), False), ("DoClip", sb.DoClip(), False),
        ("Resample 16k -> 15.2k", sb.Resample(16000, 15200), False),
        ("SpeedPerturb [95, 100, 105]", sb.SpeedPerturb(16000), False),
        ("AddNoise (white, snr 0..15)", sb.AddNoise(snr_low=0, snr_high=15), True),
        ("AddBabble (3 speakers, snr 0..15)", sb.AddBabble(snr_low=0, snr_high=15), True),
        ("TimeDomainSpecAugment", sb.TimeDomainSpecAugment(), True)]
for name, m, with_len in mods:
    us = timed((lambda: m(wav, lens)) if with_len else (lambda: m(wav)))
    print("%-34s %8.1f us per call   (read + write once at the HBM peak: %.1f us)" % (name, us, BYTES / 6553.3e3))
