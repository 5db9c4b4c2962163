"""Learnable filterbanks and their backward pass (reference:
speechbrain/processing/features.py:476-482, 512-522, 548, 691-712; lobes/features.py requires_grad).
Fixtures: outputs and GRADIENTS of the unmodified reference modules on CPU
(tests/make_golden_learnable.py)."""
import os
import sys

import numpy as np
import pytest
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import make_golden_learnable as gl  # noqa: E402

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "learnable_golden.npz")


@pytest.fixture(scope="module")
def golden():
    return np.load(GOLD)


def test_parameters_and_state_dict_match_the_reference_layout():
    import speechbrain_b200 as sb

    fb = sb.Filterbank(freeze=False, n_mels=40, param_change_factor=0.5)
    assert sorted(n for n, _ in fb.named_parameters()) == ["band", "f_central"]
    frozen = sb.Filterbank(n_mels=40)
    # stored scaled by sample_rate * param_change_factor like the reference (:476-482)
    assert torch.allclose(fb.f_central.detach() * 16000 * 0.5, frozen.f_central)
    assert sorted(n for n, _ in sb.Fbank(requires_grad=True).named_parameters()) == [
        "compute_fbanks.band", "compute_fbanks.f_central"]
    assert list(sb.Fbank().parameters()) == []


def _rel(a, b):
    return float(np.abs(a - b).max() / max(1e-30, np.abs(b).max()))


@pytest.mark.gpu
@pytest.mark.parametrize("name", list(gl.CASES))
def test_learnable_filterbank_forward_and_gradients_match_the_reference(golden, name):
    import speechbrain_b200 as sb

    kw, (s, B, T) = gl.CASES[name]
    fb = sb.Filterbank(freeze=False, **kw).cuda()
    x = gl.spec_for(s, B, T).cuda().requires_grad_(True)
    y = fb(x)
    assert np.abs(y.detach().cpu().numpy() - golden[name + "/y"]).max() < 2e-4
    (y * gl.weights_for(s, y.shape).cuda()).sum().backward()
    # gradients are sums over thousands of terms of either sign: relative to their largest entry
    assert _rel(x.grad.cpu().numpy(), golden[name + "/g_spec"]) < 2e-4
    assert _rel(fb.f_central.grad.cpu().numpy(), golden[name + "/g_fc"]) < 2e-3
    assert _rel(fb.band.grad.cpu().numpy(), golden[name + "/g_band"]) < 2e-3


@pytest.mark.gpu
def test_fbank_lobe_with_requires_grad(golden):
    import speechbrain_b200 as sb

    g = torch.Generator().manual_seed(77)
    wav = (torch.randn(2, 4000, generator=g) * 0.1).cuda()
    fbk = sb.Fbank(requires_grad=True, n_mels=40).cuda()
    y = fbk(wav)
    assert np.abs(y.detach().cpu().numpy() - golden["lobe/y"]).max() < 1e-3
    (y * gl.weights_for(77, y.shape).cuda()).sum().backward()
    assert _rel(fbk.compute_fbanks.f_central.grad.cpu().numpy(), golden["lobe/g_fc"]) < 5e-3
    assert _rel(fbk.compute_fbanks.band.grad.cpu().numpy(), golden["lobe/g_band"]) < 5e-3
    # an optimiser step moves the filters, and the next forward uses the moved ones
    opt = torch.optim.SGD(fbk.parameters(), lr=1e-12)
    before = fbk.compute_fbanks.f_central.detach().clone()
    opt.step()
    assert not torch.equal(before, fbk.compute_fbanks.f_central.detach())
    assert torch.isfinite(fbk(wav)).all()
    with pytest.raises(NotImplementedError):
        sb.MFCC(requires_grad=True)(wav)
