"""Time-domain augmentation (second half): DropChunk, DropFreq, DoClip,
Resample, SpeedPerturb and the TimeDomainSpecAugment lobe.

CPU: the numpy oracle (oracle/timedomain_np.py), driven by the product's host-side draws under
the reference's seeds, reproduces the outputs of the UNMODIFIED reference modules
(tests/golden/timedomain_golden.npz) - this pins the oracle AND the draw logic.
GPU: the kernels against the same goldens through the modules, and against the oracle at
BASELINE-sized batches.
"""
import os
import sys

import numpy as np
import pytest
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import make_golden_timedomain as gt  # noqa: E402

from oracle import timedomain_np as otd  # noqa: E402

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "timedomain_golden.npz")


@pytest.fixture(scope="module")
def golden():
    return np.load(GOLD)


@pytest.fixture(scope="module")
def sb():
    import speechbrain_b200

    return speechbrain_b200


def _oracle_run(sb, cls, kw, x, lens, seed):
    """The module's forward with the kernels replaced by the numpy oracle (host draws as is)."""
    x = np.asarray(x, dtype=np.float32)
    B, L = x.shape
    torch.manual_seed(seed)
    if cls == "TimeDomainSpecAugment":
        mod = sb.TimeDomainSpecAugment(**kw)
        y = _oracle_step(mod.speed_perturb, x, lens)
        y = _oracle_step(mod.drop_freq, y, lens)
        return _oracle_step(mod.drop_chunk, y, lens)
    return _oracle_step(getattr(sb, cls)(**kw), x, lens)


def _oracle_step(mod, x, lens):
    name = type(mod).__name__
    B, L = x.shape
    if name == "DropChunk":
        if torch.rand(1) > mod.drop_prob:
            return x.copy()
        h_valid = (torch.as_tensor(lens) * L).long().tolist()
        return otd.drop_chunks(x, mod._draw_chunks(h_valid))
    if name == "DropFreq":
        if torch.rand(1) > mod.drop_prob:
            return x.copy()
        taps, pad = mod._draw_filter()
        return otd.fir_same(x, taps.numpy(), pad).astype(np.float32)
    if name == "DoClip":
        if torch.rand(1) > mod.clip_prob:
            return x.copy()
        v = torch.rand(1)[0] * (mod.clip_high - mod.clip_low) + mod.clip_low
        return otd.clip(x, np.float32(v))
    if name == "SpeedPerturb":
        if torch.rand(1) > mod.perturb_prob:
            return x.copy()
        mod.samp_index = torch.randint(len(mod.speeds), (1,))[0]
        return _oracle_step(mod.resamplers[mod.samp_index], x, lens)
    if name == "Resample":
        if not hasattr(mod, "first_indices"):
            mod._indices_and_weights()
        if mod.orig_freq == mod.new_freq:
            return x
        return otd.resample(x, mod.weights.numpy(), mod.first_indices.numpy(), mod.conv_stride,
                            mod._output_samples(L)).astype(np.float32)
    raise AssertionError(name)


TOL = {"DropChunk": 0.0, "DoClip": 0.0, "DropFreq": 2e-6, "Resample": 2e-6, "SpeedPerturb": 2e-6,
       "TimeDomainSpecAugment": 4e-6}


@pytest.mark.parametrize("name", list(gt.CASES))
def test_oracle_and_host_draws_reproduce_the_reference(sb, golden, name):
    cls, kw, (s, B, L), with_len, seed = gt.CASES[name]
    x = gt.wav(s, B, L).numpy()
    got = _oracle_run(sb, cls, kw, x, gt.lengths_for(B).numpy(), seed)
    ref = golden[name + "/out"]
    assert got.shape == ref.shape
    assert np.abs(got - ref).max() <= TOL[cls], np.abs(got - ref).max()


def test_constructors_match_the_reference_signatures(sb):
    import inspect

    import make_golden as mg

    mg.load_reference()
    import speechbrain.lobes.augment as ref_lobes
    import speechbrain.processing.speech_augmentation as ref_sa

    for cls in ("DropChunk", "DropFreq", "DoClip", "Resample", "SpeedPerturb"):
        assert inspect.signature(getattr(sb, cls).__init__) == inspect.signature(
            getattr(ref_sa, cls).__init__), cls
    assert inspect.signature(sb.TimeDomainSpecAugment.__init__) == inspect.signature(
        ref_lobes.TimeDomainSpecAugment.__init__)


# ---------------------------------------------------------------------------------------- GPU
@pytest.mark.gpu
@pytest.mark.parametrize("name", list(gt.CASES))
def test_kernels_reproduce_the_reference(sb, golden, name):
    cls, kw, (s, B, L), with_len, seed = gt.CASES[name]
    x = gt.wav(s, B, L).cuda()
    mod = getattr(sb, cls)(**kw)
    torch.manual_seed(seed)
    y = mod(x.clone(), gt.lengths_for(B).cuda()) if with_len else mod(x.clone())
    ref = golden[name + "/out"]
    assert tuple(y.shape) == ref.shape
    assert np.abs(y.cpu().numpy() - ref).max() <= TOL[cls]
    assert y.data_ptr() != x.data_ptr()


@pytest.mark.gpu
def test_kernels_match_the_oracle_at_the_cfg2_batch(sb):
    g = torch.Generator().manual_seed(3)
    x = torch.randn(32, 256000, generator=g) * 0.1
    lens = torch.linspace(0.6, 1.0, 32)
    xd = x.cuda()
    for cls, kw, seed in (("DropChunk", {}, 1), ("DropFreq", {"drop_count_low": 2}, 2),
                          ("DoClip", {"clip_low": 0.05, "clip_high": 0.1}, 3),
                          ("Resample", {"orig_freq": 16000, "new_freq": 15200}, 4),
                          ("TimeDomainSpecAugment", {}, 5)):
        torch.manual_seed(seed)
        mod = getattr(sb, cls)(**kw)
        y = mod(xd, lens.cuda()) if cls in ("DropChunk", "TimeDomainSpecAugment") else mod(xd)
        want = _oracle_run(sb, cls, kw, x.numpy(), lens.numpy(), seed)
        assert tuple(y.shape) == want.shape
        assert np.abs(y.cpu().numpy() - want).max() <= 4e-6, cls


@pytest.mark.gpu
def test_dropchunk_noise_mode_and_three_dim_input(sb):
    """noise_factor != 0 draws its noise on the device (like the reference: not reproducible
    across devices), so: positions as drawn, values inside +- 2 * amplitude * factor, nothing
    else touched; [B, T, 1] comes back as [B, T, 1]."""
    x = (torch.randn(6, 20000) * 0.1).cuda()
    lens = torch.linspace(0.5, 1.0, 6).cuda()
    mod = sb.DropChunk(noise_factor=0.5, drop_count_low=3, drop_count_high=3)
    torch.manual_seed(9)
    y = mod(x, lens)
    torch.manual_seed(9)
    torch.rand(1)
    rounds = mod._draw_chunks((lens * 20000).long().tolist())
    touched = torch.zeros_like(x, dtype=torch.bool)
    for r in rounds:
        for b, s, n, _ in r:
            touched[b, s:s + n] = True
    assert torch.equal(y[~touched], x[~touched])
    amp = x.abs().sum(1) / (lens * 20000).long()
    for b in range(6):
        seg = y[b][touched[b]]
        assert float(seg.abs().max()) <= float(2 * amp[b] * 0.5) * (1 + 1e-6)
        assert float(seg.std()) > 0.2 * float(amp[b])
    y3 = sb.DropChunk()(x.unsqueeze(2), lens)
    assert y3.shape == (6, 20000, 1)
    with pytest.raises(NotImplementedError):
        sb.DropFreq()(torch.zeros(2, 100, 2).cuda())


@pytest.mark.gpu
@pytest.mark.parametrize("count", [0, 5, 20])
def test_dropfreq_filter_built_on_the_device_matches_the_host_chain(sb, count):
    """0 notches (identity filter), a few (sbfe_wav_notch_taps: frequencies as launch arguments)
    and more than the 16 that call takes (host-built taps): the module against the host chain of
    the reference's ops + the numpy FIR."""
    x = gt.wav(31, 3, 6000).cuda()
    mod = sb.DropFreq(drop_count_low=count, drop_count_high=count)
    torch.manual_seed(400 + count)
    y = mod(x.clone())
    torch.manual_seed(400 + count)
    assert not torch.rand(1) > mod.drop_prob
    taps, pad = mod._draw_filter()
    want = otd.fir_same(x.cpu().numpy(), taps.numpy(), pad).astype(np.float32)
    tol = 2e-6 * max(1, count)
    assert np.abs(y.cpu().numpy() - want).max() <= tol, np.abs(y.cpu().numpy() - want).max()
