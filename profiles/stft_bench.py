"""STFT on the cfg2 batch (32 x 16 s): fused-kernel instantiations (COMPLEX / POWER) against the
Stockham kernel they replace (SBFE_STFT400_OFF=1) and torch.stft (cuFFT) on the same GPU."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import speechbrain_b200 as sb
from speechbrain_b200 import ops

torch.manual_seed(0)
wav = torch.randn(32, 256000, device="cuda") * 0.1


def timed(fn, n=30):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3


for n_fft in (400, 512):
    st = sb.STFT(sample_rate=16000, n_fft=n_fft)
    tab = st.tables()
    win = torch.hamming_window(400, device="cuda")
    rows = [("fused kernel, complex", lambda: st(wav)),
            ("fused kernel, power  ", lambda: ops.stft_forward(tab, wav, power_only=True))]
    for name, fn in rows:
        print("n_fft %d  %s %7.1f us" % (n_fft, name, timed(fn)))
    os.environ["SBFE_STFT400_OFF"] = "1"
    for name, fn in rows:
        print("n_fft %d  Stockham %s %7.1f us" % (n_fft, name.split(",")[1], timed(fn)))
    del os.environ["SBFE_STFT400_OFF"]
    print("n_fft %d  torch.stft (cuFFT), complex %7.1f us" % (
        n_fft, timed(lambda: torch.stft(wav, n_fft, 160, 400, win, center=True, pad_mode="constant",
                                        return_complex=True))))
