"""Generate ``tests/golden/learnable_golden.npz``: outputs AND gradients of the UNMODIFIED reference
``Filterbank(freeze=False)`` / ``Fbank(requires_grad=True)`` (``/root/reference``) on CPU.

    python tests/make_golden_learnable.py
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import make_golden as mg  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "learnable_golden.npz")

# name: (Filterbank kwargs, (seed, B, T))
CASES = {
    "tri40": ({"n_mels": 40}, (1, 3, 50)),
    "tri80_pcf": ({"n_mels": 80, "param_change_factor": 0.5}, (2, 2, 40)),
    "gauss23": ({"n_mels": 23, "filter_shape": "gaussian"}, (3, 2, 33)),
    "linear40": ({"n_mels": 40, "log_mel": False}, (4, 2, 20)),
    "tri40_clamped": ({"n_mels": 40, "top_db": 30.0}, (5, 3, 64)),
}


def spec_for(seed, B, T, bins=201):
    g = torch.Generator().manual_seed(seed)
    s = torch.rand(B, T, bins, generator=g) ** 3 + 1e-4
    s[-1, T // 2:] *= 1e-6          # a quiet stretch: the top_db clamp bites
    return s


def weights_for(seed, shape):
    g = torch.Generator().manual_seed(1000 + seed)
    return torch.randn(*shape, generator=g)


def main():
    mg.load_reference()
    from speechbrain.lobes.features import Fbank
    from speechbrain.processing.features import Filterbank

    out = {}
    for name, (kw, (s, B, T)) in CASES.items():
        fb = Filterbank(freeze=False, **kw)
        x = spec_for(s, B, T).requires_grad_(True)
        y = fb(x)
        (y * weights_for(s, y.shape)).sum().backward()
        out[name + "/y"] = y.detach().numpy()
        out[name + "/g_spec"] = x.grad.numpy()
        out[name + "/g_fc"] = fb.f_central.grad.numpy()
        out[name + "/g_band"] = fb.band.grad.numpy()
        print("%-16s y %s  |g_fc| %.3e  |g_band| %.3e" % (name, tuple(y.shape),
              float(fb.f_central.grad.abs().max()), float(fb.band.grad.abs().max())))
    # the lobe: Fbank(requires_grad=True) on a waveform
    g = torch.Generator().manual_seed(77)
    wav = torch.randn(2, 4000, generator=g) * 0.1
    fbk = Fbank(requires_grad=True, n_mels=40)
    y = fbk(wav)
    (y * weights_for(77, y.shape)).sum().backward()
    out["lobe/y"] = y.detach().numpy()
    out["lobe/g_fc"] = fbk.compute_fbanks.f_central.grad.numpy()
    out["lobe/g_band"] = fbk.compute_fbanks.band.grad.numpy()
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
    main()
