"""Generate ``tests/golden/istft_golden.npz`` by running the UNMODIFIED reference ``ISTFT``
(speechbrain/processing/features.py:191-324, i.e. torch.istft) on CPU in the build container.

    python tests/make_golden_istft.py

Inputs are stored with the outputs (they are reference STFT outputs or seeded random spectra)."""
import os
import sys
import warnings

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import golden_cases as gc  # noqa: E402
from make_golden import load_reference  # noqa: E402

OUT = os.path.join(gc.GOLDEN_DIR, "istft_golden.npz")

# name -> (stft kwargs or None for a random spectrum, istft kwargs, sig_length, input builder)
CASES = {
    "roundtrip_default": dict(stft={"sample_rate": 16000}, istft={"sample_rate": 16000}, wav=(3, 4000), seed=301),
    "roundtrip_siglen_short": dict(stft={"sample_rate": 16000}, istft={"sample_rate": 16000}, wav=(2, 4000),
                                   seed=302, sig_length=3500),
    "roundtrip_siglen_long": dict(stft={"sample_rate": 16000}, istft={"sample_rate": 16000}, wav=(2, 3001),
                                  seed=303, sig_length=3301),
    "roundtrip_nfft512": dict(stft={"sample_rate": 16000, "n_fft": 512}, istft={"sample_rate": 16000, "n_fft": 512},
                              wav=(2, 5000), seed=304),
    "roundtrip_8k_20ms": dict(stft={"sample_rate": 8000, "n_fft": 256, "win_length": 20, "hop_length": 8},
                              istft={"sample_rate": 8000, "win_length": 20, "hop_length": 8}, wav=(2, 3000), seed=305),
    "random_spectrum": dict(stft=None, istft={"sample_rate": 16000}, spec=(2, 21, 201, 2), seed=306),
    "normalized": dict(stft={"sample_rate": 16000, "normalized_stft": True},
                       istft={"sample_rate": 16000, "normalized_stft": True}, wav=(2, 2400), seed=307),
    "multichannel": dict(stft={"sample_rate": 16000}, istft={"sample_rate": 16000}, wav=(2, 2400, 3), seed=308),
}


def main():
    _, proc = load_reference()
    out = {}
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        for name, c in CASES.items():
            rng = np.random.Generator(np.random.PCG64(c["seed"]))
            if c["stft"] is not None:
                wav = torch.from_numpy((rng.standard_normal(c["wav"]) * 0.1).astype(np.float32))
                spec = proc.STFT(**c["stft"])(wav)
            else:
                spec = torch.from_numpy(rng.standard_normal(c["spec"]).astype(np.float32))
            y = proc.ISTFT(**c["istft"])(spec, sig_length=c.get("sig_length"))
            out[name + "/spec"] = spec.numpy()
            out[name + "/out"] = y.numpy()
            print("%-26s spec %-22s -> %s" % (name, tuple(spec.shape), tuple(y.shape)))
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
    main()
