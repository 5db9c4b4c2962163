"""Where does the fused kernel deviate most from the float64 oracle on bench.py's cfg2 batch?"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench, speechbrain_b200 as sb
from oracle import frontend_np as onp

wav_t, lens_t = bench.synth_batch(0, "cpu")
wav = wav_t.numpy()
fb = sb.Fbank(n_mels=80)
got = fb(wav_t.cuda()).cpu().numpy()
consts = (fb.compute_fbanks.f_central.numpy(), fb.compute_fbanks.band.numpy())
kw = dict(n_mels=80, mel_params=consts, window=fb.compute_STFT.window.numpy())
ref64 = onp.fbank(wav, dtype=np.float64, **kw)
ref32 = onp.fbank(wav, dtype=np.float32, **kw)
d = np.abs(got - ref64)
print("max err", d.max(), "noise", np.abs(ref32 - ref64).max())
order = np.argsort(d.ravel())[::-1][:20]
for o in order:
    b, t, m = np.unravel_index(o, d.shape)
    cut = int(round(float(lens_t[b]) * wav.shape[1]))
    print("b=%d t=%d m=%d got=%.6f ref64=%.6f ref32=%.6f err=%.2e cut_frame=%.2f utt_max=%.3f"
          % (b, t, m, got[b, t, m], ref64[b, t, m], ref32[b, t, m], d[b, t, m], cut / 160.0, ref64[b].max()))
# error distribution away from the cut
mask = np.ones_like(d, bool)
for b in range(d.shape[0]):
    cut = int(round(float(lens_t[b]) * wav.shape[1])) // 160
    mask[b, max(cut - 3, 0):cut + 4] = False
print("max err away from the cut frames:", d[mask].max(), " at the cut frames:", d[~mask].max())

# the reference's own ATen chain (torch.stft -> pow/sum -> matmul -> 10 log10 clamp) run by torch on
# this GPU (cuFFT) and on the host: their distance from the float64 anchor
def torch_chain(w, dev):
    w = w.to(dev)
    win = fb.compute_STFT.window.to(dev)
    st = torch.view_as_real(torch.stft(w, 400, 160, 400, win, True, "constant", False, True,
                                       return_complex=True)).transpose(2, 1)
    power = st.pow(2).sum(-1)
    mel = torch.matmul(power, fb.compute_fbanks.fbank_matrix().to(dev))
    x_db = 10 * torch.log10(torch.clamp(mel, min=1e-10))
    floor = x_db.amax(dim=(-2, -1)) - 80.0
    return torch.max(x_db, floor.view(-1, 1, 1)).cpu().numpy()

torch.backends.cuda.matmul.allow_tf32 = False
for dev_ in ("cuda", "cpu"):
    ref_t = torch_chain(wav_t, dev_)
    dt = np.abs(ref_t - ref64)
    print("torch chain on %-4s: max err vs float64 anchor %.3e  mean %.3e   (product: max %.3e mean %.3e)"
          % (dev_, dt.max(), dt.mean(), d.max(), d.mean()))
    print("   values with err > 1e-4: torch %d, product %d of %d" % ((dt > 1e-4).sum(), (d > 1e-4).sum(), d.size))
