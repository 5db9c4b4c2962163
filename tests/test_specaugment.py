"""SpecAugment on features (reference speechbrain/lobes/augment.py:32-201).

CPU: the numpy oracle (oracle/augment_np.py) is pinned on outputs of the unmodified reference
class (tests/golden/specaugment_golden.npz, made by tests/make_golden_augment.py).
GPU: the CUDA kernel (through ops -> C ABI) against those fixtures given the same draws, and
the drop-in module against a torch-op restatement of the reference run on the same device
with the same seed (same generator stream).

Tolerance protocol as for the frontend: with ref64 = the reference in float64,
the product may be at most as far from ref64 as 2x the reference's own float32 run is, plus
one float32 ulp of the data range.
"""
import os

import numpy as np
import pytest
import torch

import make_golden_augment as mga
from oracle import augment_np as oa

ULP = 2e-6  # inputs are N(0, 1): |x| < ~5


@pytest.fixture(scope="module")
def g():
    assert os.path.exists(mga.OUT), "run python tests/make_golden_augment.py"
    return np.load(mga.OUT)


def _draws(g, name):
    def get(k):
        return g[f"{name}/{k}"] if f"{name}/{k}" in g.files else None

    warp = get("warp")
    warp = None if warp is None else (int(warp[0]), int(warp[1]))
    fr = None if get("fpos") is None else (get("fpos"), get("flen"))
    tm = None if get("tpos") is None else (get("tpos"), get("tlen"))
    return warp, fr, tm


@pytest.mark.parametrize("name", list(mga.CASES))
def test_oracle_matches_reference(g, name):
    shape, kw, _ = mga.CASES[name]
    x = g[f"{name}/x"].reshape(-1, shape[-2], shape[-1])
    warp, fr, tm = _draws(g, name)
    args = (warp, kw.get("time_warp_mode", "bicubic"), fr, tm, kw.get("replace_with_zero", True))
    ref32, ref64 = g[f"{name}/ref32"], g[f"{name}/ref64"]
    got64 = oa.spec_augment(x, *args, dtype=np.float64).reshape(shape)
    assert np.abs(got64 - ref64).max() < 1e-12
    got32 = oa.spec_augment(x, *args).reshape(shape)
    assert got32.dtype == np.float32 and got32.shape == ref32.shape
    assert np.abs(got32 - ref64).max() <= 2 * np.abs(ref32 - ref64).max() + ULP
    # masked elements are exactly the fill value in both
    if kw.get("replace_with_zero", True):
        assert np.array_equal(got32 == 0, ref32 == 0)


# ------------------------------------------------------------------------------- GPU
def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


@pytest.mark.gpu
@pytest.mark.parametrize("name", list(mga.CASES))
def test_kernel_matches_reference_given_the_draws(g, name):
    from speechbrain_b200 import ops

    shape, kw, _ = mga.CASES[name]
    x = dev(g[f"{name}/x"]).view(-1, shape[-2], shape[-1])
    warp, fr, tm = _draws(g, name)
    fr = None if fr is None else (dev(fr[0]), dev(fr[1]))
    tm = None if tm is None else (dev(tm[0]), dev(tm[1]))
    mode = kw.get("time_warp_mode", "bicubic")
    if kw.get("replace_with_zero", True):
        y = ops.specaugment_forward(x.clone(), warp, mode, fr, tm, 0.0)
    else:  # the fill is the mean of the tensor as it is when each mask is applied
        y = ops.specaugment_forward(x.clone(), warp, mode) if warp else x.clone()
        y = ops.specaugment_forward(y, None, freq_masks=fr, fill=ops.tensor_mean(y))
        y = ops.specaugment_forward(y, None, time_masks=tm, fill=ops.tensor_mean(y))
    got = y.cpu().numpy().reshape(shape)
    ref32, ref64 = g[f"{name}/ref32"], g[f"{name}/ref64"]
    assert np.abs(got - ref64).max() <= 2 * np.abs(ref32 - ref64).max() + ULP
    if kw.get("replace_with_zero", True):
        assert np.array_equal(got == 0, ref32 == 0)
    if warp is None:  # masks only: untouched elements are bit-identical
        assert np.array_equal(got, ref32)


def _reference_on_device(x, aug):
    """lobes/augment.py:106-201 restated with torch ops (same calls, same order), so that a
    seeded run on the GPU consumes the generators exactly like the reference would there."""
    import torch.nn.functional as Fn

    shape = x.shape
    if aug.apply_time_warp:
        x4 = x.unsqueeze(1) if x.dim() == 3 else x
        time, window = x4.shape[2], aug.time_warp_window
        if time - window > window:
            c = torch.randint(window, time - window, (1,))[0]
            w = torch.randint(c - window, c + window, (1,))[0] + 1
            left = Fn.interpolate(x4[:, :, :c], (w, x4.shape[3]), mode=aug.time_warp_mode, align_corners=True)
            right = Fn.interpolate(x4[:, :, c:], (time - w, x4.shape[3]), mode=aug.time_warp_mode,
                                   align_corners=True)
            x4[:, :, :w] = left
            x4[:, :, w:] = right
        x = x4.view(*shape)
    for on, dim, n, rng in ((aug.freq_mask, 2, aug.n_freq_mask, aug.freq_mask_width),
                            (aug.time_mask, 1, aug.n_time_mask, aug.time_mask_width)):
        if not on:
            continue
        x3 = x.view(-1, x.shape[-2], x.shape[-1])
        batch, D = x3.shape[0], x3.shape[dim]
        ln = torch.randint(rng[0], rng[1], (batch, n), device=x.device).unsqueeze(2)
        pos = torch.randint(0, max(1, D - ln.max()), (batch, n), device=x.device).unsqueeze(2)
        ar = torch.arange(D, device=x.device).view(1, 1, -1)
        mask = ((pos <= ar) * (ar < (pos + ln))).any(dim=1)
        mask = mask.unsqueeze(2) if dim == 1 else mask.unsqueeze(1)
        x3.masked_fill_(mask, 0.0 if aug.replace_with_zero else x3.mean())
    return x.view(*shape)


@pytest.mark.gpu
@pytest.mark.parametrize("kw,shape", [
    ({}, (32, 1601, 80)),
    ({"replace_with_zero": False}, (6, 333, 40)),
    ({"time_warp_mode": "bilinear", "n_time_mask": 5}, (3, 2, 250, 23)),
    ({"time_warp": False}, (7, 120, 80)),
])
def test_module_consumes_the_generators_like_the_reference(kw, shape):
    import speechbrain_b200 as sb

    aug = sb.SpecAugment(**kw)
    x = torch.randn(*shape, generator=torch.Generator().manual_seed(5)).cuda()
    torch.manual_seed(1234)
    want = _reference_on_device(x.clone(), aug)
    state_cpu, state_gpu = torch.get_rng_state(), torch.cuda.get_rng_state()
    torch.manual_seed(1234)
    inp = x.clone()
    got = aug(inp)
    assert got.data_ptr() == inp.data_ptr() and got.shape == x.shape   # in place, like the reference
    # both generators advanced exactly as the reference's calls advance them
    assert torch.equal(torch.get_rng_state(), state_cpu)
    assert torch.equal(torch.cuda.get_rng_state(), state_gpu)
    assert torch.equal(got == 0, want == 0) or not kw.get("replace_with_zero", True)
    torch.testing.assert_close(got, want, rtol=0, atol=2e-4)   # fp32 resampling, |x| < ~5
    changed = (got != x).float().mean().item()
    assert changed > 0.01


@pytest.mark.gpu
def test_specaugment_errors_and_trace():
    import speechbrain_b200 as sb
    from speechbrain_b200 import ops

    aug = sb.SpecAugment()
    with pytest.raises(RuntimeError):
        aug(torch.randn(2, 50, 8))                       # CPU tensor: no fallback
    with pytest.raises(AssertionError):
        sb.SpecAugment(time_warp=False, freq_mask=False, time_mask=False)
    x = torch.randn(2, 50, 8, device="cuda")
    with pytest.raises(NotImplementedError):
        ops.specaugment_forward(x, (20, 22), "nearest")
    with pytest.raises(RuntimeError):                    # a warp is not an in-place operation
        from speechbrain_b200._lib import check, lib
        check(lib().sbfe_specaugment_forward(x.data_ptr(), 2, 50, 8, 20, 22, 0, None, None, 0, None,
                                             None, 0, None, 0.0, x.data_ptr(), None))
    torch.manual_seed(3)
    traced = torch.jit.trace(aug, (x.clone(),), check_trace=False)   # the draws differ per call
    assert "sbfe::module_call" in str(traced.graph)
