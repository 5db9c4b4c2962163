"""Generate ``tests/golden/stft_modes_golden.npz``: the UNMODIFIED reference ``STFT``
(``/root/reference``) with center=False, pad_mode="reflect" and onesided=False.

    python tests/make_golden_stft_modes.py
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import make_golden as mg  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "stft_modes_golden.npz")

# name: (kwargs, (seed, B, L))
CASES = {
    "nocenter_400": ({"center": False}, (1, 2, 4000)),
    "nocenter_512_exact": ({"center": False, "n_fft": 512}, (2, 2, 512 + 160 * 3)),
    "reflect_400": ({"pad_mode": "reflect"}, (3, 3, 3217)),
    "reflect_512": ({"pad_mode": "reflect", "n_fft": 512}, (4, 2, 2000)),
    "twosided_400": ({"onesided": False}, (5, 2, 1600)),
    "twosided_512_nocenter": ({"onesided": False, "center": False, "n_fft": 512}, (6, 2, 2048)),
    "all_three_336": ({"onesided": False, "center": True, "pad_mode": "reflect", "n_fft": 336,
                       "win_length": 20}, (7, 2, 1500)),
    "reflect_multichannel": ({"pad_mode": "reflect"}, (8, 2, 2400, 2)),
}


def wav(seed, *shape):
    g = torch.Generator().manual_seed(seed)
    return torch.randn(*shape, generator=g) * 0.1


def main():
    mg.load_reference()
    from speechbrain.processing.features import STFT

    out = {}
    for name, (kw, (s, *shape)) in CASES.items():
        y = STFT(sample_rate=16000, **kw)(wav(s, *shape))
        out[name] = y.numpy()
        print("%-26s %s -> %s" % (name, tuple(shape), tuple(y.shape)))
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
    main()
