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


@pytest.mark.gpu
@pytest.mark.parametrize("B,L", [(3, 16000 + 53), (2, 5119), (1, 160 * 33), (2, 160), (32, 64000)])
@pytest.mark.parametrize("normalized", [False, True])
@pytest.mark.parametrize("n_fft", [400, 512])
def test_power_spectrum_of_the_default_mode_uses_the_fused_transform(B, L, normalized, n_fft):
    """power_only STFT at n_fft 400 / 512, hop 160 (default mode) runs the fused kernels'
    transforms (csrc/fbank400_v3.cu, fbank512.cu: POWER); same numbers as the Stockham kernel it replaces and as
    torch.stft in float64 (features.py:133-188 + spectral_magnitude power=1, :327-356).
    Tolerance: 3e-6 of the frame's largest bin (fp32 FFT of 400 points)."""
    import speechbrain_b200 as sb
    from speechbrain_b200 import ops

    torch.manual_seed(L)
    wav = torch.randn(B, L, device="cuda") * 0.1
    wav[0, L // 2:] = 0.0                                  # a zero tail: all-zero tiles
    st = sb.STFT(sample_rate=16000, n_fft=n_fft, normalized_stft=normalized)
    tab = st.tables()
    got = ops.stft_forward(tab, wav, power_only=True)
    os.environ["SBFE_STFT400_OFF"] = "1"
    try:
        old = ops.stft_forward(tab, wav, power_only=True)
    finally:
        del os.environ["SBFE_STFT400_OFF"]
    win = torch.hamming_window(400, device="cuda", dtype=torch.float64)
    ref = torch.stft(wav.double(), n_fft, 160, 400, win, center=True, pad_mode="constant",
                     normalized=normalized, onesided=True, return_complex=True)
    ref = (ref.real ** 2 + ref.imag ** 2).transpose(1, 2)
    assert got.shape == old.shape == ref.shape == (B, 1 + L // 160, n_fft // 2 + 1)
    scale = ref.amax(dim=2, keepdim=True).clamp(min=1e-30)
    err_new = float(((got.double() - ref).abs() / scale).max())
    err_old = float(((old.double() - ref).abs() / scale).max())
    assert err_new <= 3e-6, (err_new, err_old)
    assert err_old <= 3e-6, (err_new, err_old)
    assert bool((got[0, L // 320 + 4:] == 0).all())        # zeros in, exact zeros out


@pytest.mark.gpu
@pytest.mark.parametrize("B,L", [(3, 16000 + 53), (2, 5119), (1, 160 * 33), (2, 160), (32, 64000)])
@pytest.mark.parametrize("normalized", [False, True])
@pytest.mark.parametrize("n_fft", [400, 512])
def test_complex_stft_of_the_default_mode_uses_the_fused_transform(B, L, normalized, n_fft):
    """STFT.forward at n_fft 400 / 512, hop 160 (default mode; features.py:133-188) = the fused kernels'
    transforms with a complex store pass (csrc/fbank400_v3.cu, fbank512.cu: COMPLEX): against torch.stft in
    float64 and the Stockham kernel it replaces.  Tolerance: 2e-6 of the frame's largest |X[k]|."""
    import speechbrain_b200 as sb

    torch.manual_seed(L + 1)
    wav = torch.randn(B, L, device="cuda") * 0.1
    wav[0, L // 2:] = 0.0
    st = sb.STFT(sample_rate=16000, n_fft=n_fft, normalized_stft=normalized)
    got = st(wav)
    os.environ["SBFE_STFT400_OFF"] = "1"
    try:
        old = st(wav)
    finally:
        del os.environ["SBFE_STFT400_OFF"]
    win = torch.hamming_window(400, device="cuda", dtype=torch.float64)
    ref = torch.stft(wav.double(), n_fft, 160, 400, win, center=True, pad_mode="constant",
                     normalized=normalized, onesided=True, return_complex=True)
    ref = torch.view_as_real(ref).transpose(1, 2)                  # [B, T, bins, 2]
    assert got.shape == old.shape == ref.shape == (B, 1 + L // 160, n_fft // 2 + 1, 2)
    scale = ref.pow(2).sum(-1).sqrt().amax(dim=2, keepdim=True).unsqueeze(-1).clamp(min=1e-30)
    err_new = float(((got.double() - ref).abs() / scale).max())
    err_old = float(((old.double() - ref).abs() / scale).max())
    assert err_new <= 2e-6, (err_new, err_old)
    assert err_old <= 2e-6, (err_new, err_old)
    assert bool((got[0, L // 320 + 4:] == 0).all())
    assert bool((got[..., 0, 1] == 0).all()) and bool((got[..., n_fft // 2, 1] == 0).all())   # real bins
