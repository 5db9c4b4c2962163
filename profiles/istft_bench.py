"""Times STFT and ISTFT on the cfg2 batch shape (32 x 16 s), eager calls, CUDA events; the reference's
torch.istft on the same GPU beside it.   python profiles/istft_bench.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import speechbrain_b200 as sb

torch.manual_seed(0)
wav = torch.randn(32, 256000, device="cuda") * 0.1
stft, istft = sb.STFT(sample_rate=16000), sb.ISTFT(sample_rate=16000)
spec = stft(wav)
win = torch.hamming_window(400, device="cuda")


def ref_istft(x):  # what the reference class executes (processing/features.py:300-318)
    z = torch.complex(x[..., 0], x[..., 1]).permute(0, 2, 1)
    return torch.istft(z, n_fft=400, hop_length=160, win_length=400, window=win, center=True, onesided=True)


for name, fn in (("sbfe STFT  ", lambda: stft(wav)), ("sbfe ISTFT ", lambda: istft(spec)),
                 ("torch.istft", lambda: ref_istft(spec))):
    for _ in range(3):
        out = fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        out = fn()
    e1.record()
    torch.cuda.synchronize()
    print("%s 32 x 16 s: %8.1f us per call" % (name, e0.elapsed_time(e1) * 100))
print("max |istft(stft(x)) - x| = %.2e (sbfe), %.2e (torch.istft on the sbfe spectrum)" % (
    float((istft(spec) - wav).abs().max()), float((ref_istft(spec) - wav).abs().max())))
