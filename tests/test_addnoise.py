"""AddNoise, white-noise mode (next row N1): oracle pinned on the unmodified reference's outputs
(tests/golden/addnoise_golden.npz, tests/make_golden_addnoise.py); kernels against the oracle."""
import os
import sys

import numpy as np
import pytest
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import make_golden_addnoise as ga  # noqa: E402

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import timedomain_np as otd  # noqa: E402

TOL = 2e-6   # float32 chain; the sum of |x| is accumulated in another order than torch.sum


@pytest.fixture(scope="module")
def golden():
    return np.load(ga.OUT)


def _replay(kw, x, lens, seed, device="cpu"):
    """The reference's draws (:131-148) from the seed, then the oracle's arithmetic."""
    torch.manual_seed(seed)
    if torch.rand(1) > kw.get("mix_prob", 1.0):
        return x.cpu().numpy().copy()
    lo, hi = kw.get("snr_low", 0), kw.get("snr_high", 0)
    snr = torch.rand(len(x), 1, device=device) * (hi - lo) + lo
    noise = torch.randn_like(x)
    return otd.add_noise(x.cpu().numpy(), noise.cpu().numpy(), lens.cpu().numpy(),
                         snr.cpu().numpy().reshape(-1), kw.get("normalize", False))


@pytest.mark.parametrize("name", list(ga.CASES))
def test_oracle_reproduces_the_reference(golden, name):
    kw, (s, B, L, gain), seed = ga.CASES[name]
    x, lens = ga.wav(s, B, L, gain)
    got = _replay(kw, x, lens, seed)
    ref = golden[name + "/out"]
    assert got.shape == ref.shape
    assert np.abs(got - ref).max() <= TOL, np.abs(got - ref).max()


def test_constructor_matches_the_reference_signature():
    import inspect

    import speechbrain_b200 as sb

    names = list(inspect.signature(sb.AddNoise.__init__).parameters)[1:]
    assert names == ["csv_file", "csv_keys", "sorting", "num_workers", "snr_low", "snr_high", "pad_noise",
                     "mix_prob", "start_index", "normalize", "replacements", "noise_sample_rate",
                     "clean_sample_rate"]
    with pytest.raises(NotImplementedError):
        sb.AddNoise(csv_file="noise.csv")


@pytest.mark.gpu
@pytest.mark.parametrize("name", list(ga.CASES))
def test_kernels_match_the_oracle_on_the_reference_cases(name):
    """Same seed, same device: the module's draws are the reference's calls, so replaying them
    in front of the oracle gives what the reference computes on this GPU."""
    import speechbrain_b200 as sb

    kw, (s, B, L, gain), seed = ga.CASES[name]
    x, lens = ga.wav(s, B, L, gain)
    x, lens = x.cuda(), lens.cuda()
    torch.manual_seed(seed)
    y = sb.AddNoise(**kw)(x.clone(), lens)
    want = _replay(kw, x, lens, seed, device="cuda")
    assert tuple(y.shape) == want.shape and y.data_ptr() != x.data_ptr()
    assert np.abs(y.cpu().numpy() - want).max() <= TOL


@pytest.mark.gpu
@pytest.mark.parametrize("normalize", [False, True])
def test_kernels_match_the_oracle_at_the_cfg2_batch(normalize):
    from speechbrain_b200 import ops

    torch.manual_seed(3)
    B, L = 32, 256000
    x = torch.randn(B, L, device="cuda") * (1.5 if normalize else 0.1)
    lens = torch.linspace(0.6, 1.0, B, device="cuda")
    snr = torch.rand(B, device="cuda") * 20 - 5
    noise = torch.randn_like(x)
    y = ops.wav_add_noise(x, noise, lens, snr, normalize)
    want = otd.add_noise(x.cpu().numpy(), noise.cpu().numpy(), lens.cpu().numpy(), snr.cpu().numpy(), normalize)
    assert np.abs(y.cpu().numpy() - want).max() <= TOL * max(1.0, float(np.abs(want).max()))
    y3 = ops.wav_add_noise(x, noise, lens, snr, normalize)
    assert torch.equal(y, y3)                                 # fixed summation order: deterministic


# ---- AddBabble -------------------------------------------------------------------------------
def _replay_babble(kw, x, lens, seed, device="cpu"):
    torch.manual_seed(seed)
    if torch.rand(1) > kw.get("mix_prob", 1):
        return x.cpu().numpy().copy()
    lo, hi = kw.get("snr_low", 0), kw.get("snr_high", 0)
    snr = torch.rand(len(x), 1, device=device) * (hi - lo) + lo
    return otd.add_babble(x.cpu().numpy(), lens.cpu().numpy(), snr.cpu().numpy().reshape(-1),
                          kw.get("speaker_count", 3))


@pytest.mark.parametrize("name", list(ga.BABBLE))
def test_babble_oracle_reproduces_the_reference(golden, name):
    kw, (s, B, L, gain), seed = ga.BABBLE[name]
    x, lens = ga.wav(s, B, L, gain)
    got = _replay_babble(kw, x, lens, seed)
    ref = golden[name + "/out"]
    assert got.shape == ref.shape
    assert np.abs(got - ref).max() <= TOL, np.abs(got - ref).max()


def test_babble_constructor_matches_the_reference_signature():
    import inspect

    import speechbrain_b200 as sb

    assert list(inspect.signature(sb.AddBabble.__init__).parameters)[1:] == [
        "speaker_count", "snr_low", "snr_high", "mix_prob"]


@pytest.mark.gpu
@pytest.mark.parametrize("name", list(ga.BABBLE))
def test_babble_kernels_match_the_oracle_on_the_reference_cases(name):
    import speechbrain_b200 as sb

    kw, (s, B, L, gain), seed = ga.BABBLE[name]
    x, lens = ga.wav(s, B, L, gain)
    x, lens = x.cuda(), lens.cuda()
    torch.manual_seed(seed)
    y = sb.AddBabble(**kw)(x.clone(), lens)
    want = _replay_babble(kw, x, lens, seed, device="cuda")
    assert tuple(y.shape) == want.shape and y.data_ptr() != x.data_ptr()
    assert np.abs(y.cpu().numpy() - want).max() <= TOL


@pytest.mark.gpu
def test_babble_kernels_match_the_oracle_at_the_cfg2_batch():
    from speechbrain_b200 import ops

    torch.manual_seed(4)
    B, L = 32, 256000
    x = torch.randn(B, L, device="cuda") * 0.1
    lens = torch.linspace(0.6, 1.0, B, device="cuda")
    snr = torch.rand(B, device="cuda") * 20 - 5
    y = ops.wav_add_babble(x, lens, snr, 3)
    want = otd.add_babble(x.cpu().numpy(), lens.cpu().numpy(), snr.cpu().numpy(), 3)
    assert np.abs(y.cpu().numpy() - want).max() <= TOL * max(1.0, float(np.abs(want).max()))
    assert torch.equal(y, ops.wav_add_babble(x, lens, snr, 3))     # deterministic
