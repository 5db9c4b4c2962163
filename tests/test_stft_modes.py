"""torch.stft's other modes in the stand-alone STFT module (missing 6):
center=False, pad_mode="reflect", onesided=False (speechbrain/processing/features.py:101-112,
161-172).  Fixtures: outputs of the unmodified reference STFT (tests/make_golden_stft_modes.py).
CPU: the numpy oracle reproduces them; GPU: the kernels (Stockham path for n_fft 400 / 512,
direct DFT for n_fft 336) do, and the lobes refuse a non-default STFT."""
import os
import sys

import numpy as np
import pytest
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import make_golden_stft_modes as gm  # noqa: E402

from oracle import frontend_np as onp  # noqa: E402

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "stft_modes_golden.npz")


@pytest.fixture(scope="module")
def golden():
    return np.load(GOLD)


@pytest.mark.parametrize("name", list(gm.CASES))
def test_oracle_reproduces_the_reference(golden, name):
    kw, (s, *shape) = gm.CASES[name]
    got = onp.stft(gm.wav(s, *shape).numpy(), **kw)
    ref = golden[name]
    assert got.shape == ref.shape
    assert np.abs(got - ref).max() <= 2e-5 * max(1.0, np.abs(ref).max())


@pytest.mark.gpu
@pytest.mark.parametrize("name", list(gm.CASES))
def test_kernels_reproduce_the_reference(golden, name):
    import speechbrain_b200 as sb

    kw, (s, *shape) = gm.CASES[name]
    got = sb.STFT(sample_rate=16000, **kw)(gm.wav(s, *shape).cuda()).cpu().numpy()
    ref = golden[name]
    assert got.shape == ref.shape
    assert np.abs(got - ref).max() <= 2e-5 * max(1.0, np.abs(ref).max())


@pytest.mark.gpu
def test_error_behaviour_of_the_modes():
    import speechbrain_b200 as sb

    x = torch.zeros(2, 300).cuda()
    with pytest.raises(RuntimeError):            # torch.stft: fewer samples than n_fft
        sb.STFT(sample_rate=16000, center=False)(x)
    with pytest.raises(RuntimeError):            # torch.stft: reflection longer than the signal
        sb.STFT(sample_rate=16000, pad_mode="reflect")(torch.zeros(1, 150).cuda())
    with pytest.raises(NotImplementedError):
        sb.STFT(sample_rate=16000, pad_mode="circular")(x)
    fb = sb.Fbank()
    fb.compute_STFT = sb.STFT(sample_rate=16000, center=False)
    with pytest.raises(NotImplementedError):
        fb(torch.zeros(2, 4000).cuda())
