"""Generate ``tests/golden/specaugment_golden.npz`` by running the UNMODIFIED reference
``speechbrain.lobes.augment.SpecAugment`` (``/root/reference``) on CPU in the build container.

    python tests/make_golden_augment.py

For every case the reference is run under ``torch.manual_seed(seed)``; the same seed is then
replayed to recover the draws the run consumed (c, w, mask lengths / positions: the exact
``torch.randint`` calls of lobes/augment.py:129-130, 177-183 in order), so the fixture holds
input, draws and the reference's output.
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import make_golden as mg  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "specaugment_golden.npz")

CASES = {
    # name: (shape, kwargs, seed)
    "default_8x120x80": ((8, 120, 80), {}, 11),
    "warp_only_3x61x23": ((3, 61, 23), {"freq_mask": False, "time_mask": False}, 12),
    "bilinear_4x200x40": ((4, 200, 40), {"time_warp_mode": "bilinear"}, 13),
    "masks_only_5x90x80": ((5, 90, 80), {"time_warp": False}, 14),
    "mean_fill_4x150x80": ((4, 150, 80), {"replace_with_zero": False}, 15),
    "wide_window_2x400x80": ((2, 400, 80), {"time_warp_window": 40, "n_freq_mask": 3,
                                           "n_time_mask": 4, "time_mask_width": (5, 30)}, 16),
    "short_no_warp_2x9x12": ((2, 9, 12), {"freq_mask_width": (0, 5), "time_mask_width": (0, 4)}, 17),
    "four_dim_2x2x70x40": ((2, 2, 70, 40), {}, 18),
}


def replay_draws(aug, shape, seed):
    """The draws of one forward pass, made by the same calls in the same order."""
    torch.manual_seed(seed)
    time, fea = shape[-2], shape[-1]
    batch = int(np.prod(shape[:-2]))
    d = {}
    if aug.apply_time_warp:
        window = aug.time_warp_window
        if time - window > window:
            c = torch.randint(window, time - window, (1,))[0]
            w = torch.randint(c - window, c + window, (1,))[0] + 1
            d["warp"] = np.array([int(c), int(w)], dtype=np.int64)
    for key, on, D, n, rng in (("f", aug.freq_mask, fea, aug.n_freq_mask, aug.freq_mask_width),
                               ("t", aug.time_mask, time, aug.n_time_mask, aug.time_mask_width)):
        if on:
            ln = torch.randint(rng[0], rng[1], (batch, n))
            pos = torch.randint(0, max(1, D - ln.max()), (batch, n))
            d[key + "len"], d[key + "pos"] = ln.numpy(), pos.numpy()
    return d


def main():
    mg.load_reference()
    import speechbrain.lobes.augment as ref_aug

    out = {}
    for name, (shape, kwargs, seed) in CASES.items():
        g = torch.Generator().manual_seed(1000 + seed)
        x = torch.randn(*shape, generator=g)
        aug = ref_aug.SpecAugment(**kwargs)
        out[name + "/x"] = x.numpy().copy()
        torch.manual_seed(seed)
        y = aug(x.clone())
        out[name + "/ref32"] = y.numpy().copy()
        for k, v in replay_draws(aug, shape, seed).items():
            out[name + "/" + k] = v
        # fp64 anchor of the same run
        torch.manual_seed(seed)
        out[name + "/ref64"] = aug(x.double().clone()).numpy().copy()
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, "%.1f kB" % (os.path.getsize(OUT) / 1e3))


if __name__ == "__main__":
    main()
