"""ISTFT (reference speechbrain/processing/features.py:191-324 = torch.istft).

CPU: the numpy oracle (oracle/frontend_np.py: istft) is pinned on outputs of the unmodified
reference class (tests/golden/istft_golden.npz, made by tests/make_golden_istft.py).
GPU: the CUDA kernels (drop-in module -> ops -> C ABI) against those fixtures, the float64 oracle,
and a size-independent property at the cfg2 batch size: ISTFT(STFT(x)) == x.
Tolerance: outputs are O(0.1 .. 1); float32 inverse DFT of 400 points -> 2e-6 absolute."""
import os

import numpy as np
import pytest
import torch

import make_golden_istft as mgi
from oracle import frontend_np as onp

TOL = 2e-6


@pytest.fixture(scope="module")
def g():
    assert os.path.exists(mgi.OUT), "run python tests/make_golden_istft.py"
    return np.load(mgi.OUT)


def _kw(case):
    return {k: v for k, v in case["istft"].items() if k != "normalized_stft"}


@pytest.mark.parametrize("name", sorted(mgi.CASES))
def test_oracle_matches_reference(g, name):
    case = mgi.CASES[name]
    want = g[name + "/out"]
    for dt in (np.float32, np.float64):
        got = onp.istft(g[name + "/spec"], sig_length=case.get("sig_length"), dtype=dt, **_kw(case))
        assert got.shape == want.shape
        assert float(np.abs(got - want).max()) <= TOL * max(1.0, float(np.abs(want).max()))


def test_oracle_refuses_vanishing_envelope():
    spec = np.zeros((1, 3, 201, 2), np.float32)
    with pytest.raises(RuntimeError):   # hop > window: gaps between the frames
        onp.istft(spec, win_length=5, hop_length=10, n_fft=400)


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(mgi.CASES))
def test_cuda_matches_reference(g, name):
    import speechbrain_b200 as sb

    case = mgi.CASES[name]
    spec = torch.from_numpy(g[name + "/spec"]).cuda()
    mod = sb.ISTFT(**case["istft"])
    got = mod(spec, sig_length=case.get("sig_length")).cpu().numpy()
    want = g[name + "/out"]
    ref64 = onp.istft(g[name + "/spec"], sig_length=case.get("sig_length"), dtype=np.float64, **_kw(case))
    assert got.shape == want.shape
    scale = max(1.0, float(np.abs(want).max()))
    assert float(np.abs(got - want).max()) <= TOL * scale, float(np.abs(got - want).max())
    # no less accurate than the reference's own float32 run (SURVEY.md 8c protocol)
    assert float(np.abs(got - ref64).max()) <= 2 * float(np.abs(want - ref64).max()) + 2e-7 * scale


@pytest.mark.gpu
def test_cuda_roundtrip_at_full_size_and_errors():
    import speechbrain_b200 as sb

    torch.manual_seed(5)
    wav = torch.randn(32, 256000, device="cuda") * 0.1          # the cfg2 batch
    spec = sb.STFT(sample_rate=16000)(wav)
    back = sb.ISTFT(sample_rate=16000)(spec)
    assert back.shape == wav.shape
    assert float((back - wav).abs().max()) < 2e-6
    assert sb.ISTFT(sample_rate=16000)(spec, sig_length=1000).shape == (32, 1000)
    with pytest.raises(RuntimeError):                            # same refusal as torch.istft
        sb.ISTFT(sample_rate=16000, win_length=5, hop_length=10, n_fft=400)(spec[:1, :4])
    with pytest.raises(RuntimeError):                            # no CPU fallback
        sb.ISTFT(sample_rate=16000)(spec.cpu())
    with pytest.raises(NotImplementedError):
        sb.ISTFT(sample_rate=16000, onesided=False)(spec)


@pytest.mark.gpu
def test_cuda_direct_dft_paths_for_other_n_fft():
    """n_fft / 2 with a prime factor other than 2 and 5 (here 336 = 2^4 * 21) takes the direct-DFT
    kernels, forward and inverse: checked against the float64 oracle."""
    import speechbrain_b200 as sb

    rng = np.random.default_rng(41)
    wav = (rng.standard_normal((2, 3000)) * 0.1).astype(np.float32)
    kw = dict(sample_rate=16000, n_fft=336, win_length=20, hop_length=10)
    spec = sb.STFT(**kw)(torch.from_numpy(wav).cuda())
    want = onp.stft(wav, dtype=np.float64, **kw)
    assert spec.shape == want.shape
    assert float(np.abs(spec.cpu().numpy() - want).max()) < 2e-5      # |X| up to ~10
    x = rng.standard_normal((2, 9, 169, 2)).astype(np.float32)
    got = sb.ISTFT(**kw)(torch.from_numpy(x).cuda()).cpu().numpy()
    ref = onp.istft(x, dtype=np.float64, **kw)
    assert got.shape == ref.shape
    assert float(np.abs(got - ref).max()) < 2e-6 * max(1.0, float(np.abs(ref).max()))
    back = sb.ISTFT(**kw)(spec).cpu().numpy()
    assert float(np.abs(back - wav[:, :back.shape[1]]).max()) < 2e-6
