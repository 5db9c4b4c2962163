"""Pins the numpy oracle against outputs of the unmodified reference (golden
fixtures made by tests/make_golden.py) and against the known-answer vectors of
the reference's own tests/unittests/test_features.py.  CPU only."""
import numpy as np
import pytest

import golden_cases as gc
from oracle import frontend_np as onp

LOBE_CASES = [n for n, s in gc.CASES.items() if s["kind"] in ("fbank", "mfcc", "fbank+norm")]
OP_CASES = [n for n, s in gc.CASES.items()
            if s["kind"] in ("stft", "magnitude", "filterbank", "dct", "deltas", "context")]
SEQ_CASES = [n for n, s in gc.CASES.items() if s["kind"] == "norm_seq"]


def _lobe(spec, x, dtype, consts=None):
    kw = dict(spec["kwargs"])
    if consts is not None:
        kw["mel_params"], kw["window"] = (consts[0], consts[1]), consts[2]
    if "deltas" in kw:
        kw["deltas_"] = kw.pop("deltas")
    fn = onp.mfcc if spec["kind"] == "mfcc" else onp.fbank
    return fn(x, dtype=dtype, **kw)


@pytest.mark.parametrize("name", LOBE_CASES)
def test_lobe_matches_reference(golden, name):
    spec = gc.CASES[name]
    x = spec["make_input"]()
    np.testing.assert_allclose(gc.input_checksum(x), golden[f"{name}/in_ck"], rtol=1e-12)
    ref32 = golden[f"{name}/ref32"]
    # tight checks use the reference's own float32 filter/window constants ...
    consts = tuple(golden[f"{name}/{k}"] for k in ("fc", "band", "window"))
    got32 = _lobe(spec, x, np.float32, consts)
    assert got32.shape == ref32.shape and got32.dtype == np.float32
    # ... the oracle's numpy-built constants differ from torch's in the last ulp,
    # which the 10**m - 1 cancellation turns into <~3e-4 dB on the lowest filters
    own32 = _lobe(spec, x, np.float32)
    assert np.abs(own32 - got32).max() < 5e-4 * (11 if spec["kwargs"].get("context") or spec["kind"] == "mfcc" else 1)
    has64 = f"{name}/ref64" in golden.files
    if has64:
        ref64 = golden[f"{name}/ref64"]
        got64 = _lobe(spec, x, np.float64, consts)
        # fp64 restatement vs fp64 run of the reference: only constant-rounding noise
        assert np.abs(got64 - ref64).max() < 2e-5
        noise = np.abs(ref32 - ref64).max()
        # the fp32 oracle is as close to the fp64 anchor as the reference itself is
        assert np.abs(got32 - ref64).max() <= 2 * noise + 1e-5
    assert np.abs(got32 - ref32).max() < 2e-3
    assert np.abs(got32 - ref32).mean() < 2e-5
    if spec["kind"] == "fbank+norm":
        norm = onp.InputNormalizationOracle(**spec["norm"])
        out = norm(got32.copy(), spec["lengths"])
        assert np.abs(out - golden[f"{name}/norm32"]).max() < 2e-3
        if has64:
            norm = onp.InputNormalizationOracle(**spec["norm"])
            out64 = norm(_lobe(spec, x, np.float64, consts), spec["lengths"])
            assert np.abs(out64 - golden[f"{name}/norm64"]).max() < 2e-5


@pytest.mark.parametrize("name", OP_CASES)
def test_op_matches_reference(golden, name):
    spec = gc.CASES[name]
    x = spec["make_input"]()
    np.testing.assert_allclose(gc.input_checksum(x), golden[f"{name}/in_ck"], rtol=1e-12)
    ref = golden[f"{name}/ref32"]
    kind, kw = spec["kind"], dict(spec["kwargs"])
    if kind == "stft":
        got = onp.stft(x, **kw)
        tol = 2e-4
    elif kind == "magnitude":
        got = onp.spectral_magnitude(x, **kw)
        tol = 1e-5
    elif kind == "filterbank":
        got = onp.filterbank(x, **kw)
        tol = 1e-4
    elif kind == "dct":
        kw.pop("input_size")
        got = onp.dct(x, **kw)
        tol = 1e-4
    elif kind == "deltas":
        kw.pop("input_size")
        got = onp.deltas(x, **kw)
        tol = 1e-5
    else:
        got = onp.context_window(x, **kw)
        tol = 0.0
    assert got.shape == ref.shape
    assert np.abs(got - ref).max() <= tol * max(1.0, np.abs(ref).max())


@pytest.mark.parametrize("name", SEQ_CASES)
def test_norm_sequence_matches_reference(golden, name):
    spec = gc.CASES[name]
    norm = onp.InputNormalizationOracle(**spec["norm"])
    for i, st in enumerate(spec["steps"]):
        x = gc.synth_feats(st["seed"], spec["shape"])
        norm.training = st["training"]
        y = norm(x.copy(), st["lengths"], spk_ids=st["spk_ids"], epoch=st["epoch"])
        ref = golden[f"{name}/step{i}"]
        assert y.shape == ref.shape
        assert np.abs(y - ref).max() < 5e-5 * max(1.0, np.abs(ref).max())
        if spec["norm"]["norm_type"] == "global":
            assert norm.count == int(golden[f"{name}/step{i}_count"])
            np.testing.assert_allclose(np.broadcast_to(norm.glob_mean, golden[f"{name}/step{i}_glob_mean"].shape),
                                       golden[f"{name}/step{i}_glob_mean"], rtol=1e-5, atol=1e-5)


# ---- known-answer vectors of reference tests/unittests/test_features.py -----
def test_ka_deltas_of_ones_is_zero():  # test_features.py:4-14
    assert np.all(onp.deltas(np.ones((10, 101, 20), np.float32)) == 0)


def test_ka_context_window():  # test_features.py:17-39
    x = np.array([1, 2, 3], np.float32)[None, :, None]
    want = np.array([[0, 1, 2], [1, 2, 3], [2, 3, 0]], np.float32)[None]
    assert np.array_equal(onp.context_window(x, 1, 1), want)
    x = np.random.rand(2, 10, 5).astype(np.float32)
    assert np.array_equal(onp.context_window(x, 0, 0), x)


def test_ka_filterbank_floor_and_batch_independence():  # test_features.py:60-88
    fb = onp.filterbank(np.zeros((10, 101, 201), np.float32))
    assert np.array_equal(fb, np.full_like(fb, -100.0))
    assert onp.amplitude_to_db(np.zeros((1, 1, 1), np.float32))[0, 0, 0] == -100.0
    rng = np.random.default_rng(0)
    a = (rng.random((1, 101, 201)) * 10).astype(np.float32)
    b = rng.random((1, 101, 201)).astype(np.float32)
    both = onp.filterbank(np.concatenate([a, b]))
    assert np.abs(onp.filterbank(a)[0] - both[0]).sum() < 8e-5
    assert np.abs(onp.filterbank(b)[0] - both[1]).sum() < 8e-5


def test_ka_input_normalization():  # test_features.py:100-119
    x = np.array([1, 2, 3, 0, 0, 0], np.float32)[None, :, None]
    out = onp.InputNormalizationOracle()(x, [0.5]).squeeze()
    assert np.array_equal(out, np.array([-1, 0, 1, -2, -2, -2], np.float32))


def test_ka_multimic_filterbank():  # test_features.py:122-132
    x = np.random.default_rng(1).random((10, 101, 201)).astype(np.float32)
    one = onp.filterbank(x)
    two = onp.filterbank(np.stack([x, x], -1))[..., 0]
    assert np.abs(one - two).sum() < 5e-5


def test_ka_doctest_shapes():  # processing/features.py:95-98, lobes/features.py:75-79,205-209
    x = np.random.default_rng(2).standard_normal((10, 16000)).astype(np.float32)
    assert onp.stft(x).shape == (10, 101, 201, 2)
    assert onp.fbank(x).shape == (10, 101, 40)
    assert onp.mfcc(x).shape == (10, 101, 660)
    assert onp.spectral_magnitude(np.array([[3.0, 4.0]], np.float32), power=0.5)[0] == pytest.approx(5.0)
    with pytest.raises(ValueError):
        onp.dct_matrix(10, 20)
