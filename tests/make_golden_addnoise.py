"""Generate ``tests/golden/addnoise_golden.npz``: the UNMODIFIED reference ``AddNoise`` in its
white-noise mode and ``AddBabble`` (``/root/reference``, speechbrain/processing/speech_augmentation.py:30-168,
824-906) on CPU
under ``torch.manual_seed(seed)``.

    python tests/make_golden_addnoise.py

The fixture holds the reference's output; input, lengths and the draws (``torch.rand(1)``,
``torch.rand(B, 1)``, ``torch.randn_like``) are replayed from the seeds by the tests.
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import make_golden as mg  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "addnoise_golden.npz")

# name: (kwargs, input (seed, B, L, gain), run seed)
CASES = {
    "snr0": ({}, (21, 4, 16000, 0.1), 201),
    "snr_range": ({"snr_low": 5, "snr_high": 25}, (22, 3, 8003, 0.1), 202),
    "normalize_loud": ({"snr_low": -10, "snr_high": 0, "normalize": True}, (23, 5, 4000, 2.0), 203),
    "normalize_quiet": ({"snr_low": 10, "snr_high": 20, "normalize": True}, (24, 2, 1777, 0.05), 204),
    "skipped": ({"mix_prob": 0.0}, (25, 2, 1000, 0.1), 205),
}


# AddBabble: name: (kwargs, input (seed, B, L, gain), run seed)
BABBLE = {
    "babble_default": ({}, (31, 5, 8000, 0.1), 301),
    "babble_snr_range": ({"speaker_count": 2, "snr_low": 5, "snr_high": 20}, (32, 4, 5003, 0.2), 302),
    "babble_wraps": ({"speaker_count": 5, "snr_low": -5, "snr_high": 5}, (33, 3, 3000, 0.1), 303),
    "babble_skipped": ({"mix_prob": 0.0}, (34, 2, 1000, 0.1), 304),
}


def wav(seed, B, L, gain):
    g = torch.Generator().manual_seed(seed)
    x = torch.randn(B, L, generator=g) * gain
    lens = torch.linspace(0.55, 1.0, B)
    for b in range(B):
        x[b, int(round(float(lens[b]) * L)):] = 0.0
    return x, lens


def main():
    mg.load_reference()
    import speechbrain.processing.speech_augmentation as ref_sa

    out = {}
    for name, (kw, (s, B, L, gain), seed) in CASES.items():
        x, lens = wav(s, B, L, gain)
        mod = ref_sa.AddNoise(**kw)
        torch.manual_seed(seed)
        y = mod(x.clone(), lens)
        out[name + "/out"] = y.numpy()
        print("%-18s %s max |y| %.3f" % (name, tuple(y.shape), float(y.abs().max())))
    for name, (kw, (s, B, L, gain), seed) in BABBLE.items():
        x, lens = wav(s, B, L, gain)
        mod = ref_sa.AddBabble(**kw)
        torch.manual_seed(seed)
        y = mod(x.clone(), lens)
        out[name + "/out"] = y.numpy()
        print("%-18s %s max |y| %.3f" % (name, tuple(y.shape), float(y.abs().max())))
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
    main()
