"""Generate ``tests/golden/timedomain_golden.npz`` by running the UNMODIFIED reference
``speechbrain.processing.speech_augmentation`` / ``speechbrain.lobes.augment`` modules
(``/root/reference``) on CPU in the build container.

    python tests/make_golden_timedomain.py

Every case runs the reference under ``torch.manual_seed(seed)``; the product replays the same seed
(the draws are the same ``torch.rand`` / ``torch.randint`` calls on the default generator), so the
fixture holds input, lengths and the reference's output.
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import make_golden as mg  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "timedomain_golden.npz")


def wav(seed, B, L):
    g = torch.Generator().manual_seed(seed)
    return torch.randn(B, L, generator=g) * 0.1


# name: (class name, kwargs, input (seed, B, L), needs lengths, run seed)
CASES = {
    "dropchunk_default": ("DropChunk", {}, (1, 4, 16000), True, 101),
    "dropchunk_window": ("DropChunk", {"drop_start": 100, "drop_end": 2000, "drop_length_low": 50,
                                       "drop_length_high": 400, "drop_count_low": 2,
                                       "drop_count_high": 6}, (2, 3, 4000), True, 102),
    "dropchunk_negative_bounds": ("DropChunk", {"drop_start": -3000, "drop_end": -100,
                                                "drop_length_low": 10, "drop_length_high": 200},
                                  (3, 5, 8000), True, 103),
    "dropchunk_skipped": ("DropChunk", {"drop_prob": 0.0}, (4, 2, 1000), True, 104),
    "dropfreq_default": ("DropFreq", {}, (5, 3, 8000), False, 105),
    "dropfreq_three": ("DropFreq", {"drop_count_low": 3, "drop_count_high": 3, "drop_width": 0.1},
                       (6, 2, 5003), False, 106),
    "dropfreq_short": ("DropFreq", {}, (7, 2, 60), False, 107),
    "doclip_default": ("DoClip", {"clip_low": 0.05, "clip_high": 0.2}, (8, 3, 4001), False, 108),
    "resample_16k_8k": ("Resample", {"orig_freq": 16000, "new_freq": 8000}, (9, 2, 5217), False, 109),
    "resample_16k_15200": ("Resample", {"orig_freq": 16000, "new_freq": 15200}, (10, 3, 16000), False, 110),
    "resample_16k_16800": ("Resample", {"orig_freq": 16000, "new_freq": 16800}, (11, 2, 9999), False, 111),
    "resample_8k_16k": ("Resample", {"orig_freq": 8000, "new_freq": 16000}, (12, 2, 3000), False, 112),
    "speedperturb_default": ("SpeedPerturb", {"orig_freq": 16000}, (13, 2, 8000), False, 113),
    "speedperturb_95_105": ("SpeedPerturb", {"orig_freq": 16000, "speeds": [95, 100, 105]},
                            (14, 2, 8000), False, 115),
    "tdsa_default": ("TimeDomainSpecAugment", {}, (15, 4, 16000), True, 116),
    "tdsa_speed80": ("TimeDomainSpecAugment", {"speeds": [80], "drop_chunk_length_low": 200,
                                               "drop_chunk_length_high": 900}, (16, 3, 12000), True, 117),
}


def lengths_for(B):
    return torch.linspace(0.55, 1.0, B)


def main():
    mg.load_reference()
    import speechbrain.lobes.augment as ref_lobes
    import speechbrain.processing.speech_augmentation as ref_sa

    out = {}
    for name, (cls, kw, (s, B, L), with_len, seed) in CASES.items():
        x = wav(s, B, L)
        mod = getattr(ref_lobes if cls == "TimeDomainSpecAugment" else ref_sa, cls)(**kw)
        torch.manual_seed(seed)
        y = mod(x.clone(), lengths_for(B)) if with_len else mod(x.clone())
        out[name + "/out"] = y.numpy()
        print("%-28s %s -> %s" % (name, tuple(x.shape), tuple(y.shape)))
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
    main()
