"""GPU parity: the CUDA product (through the drop-in modules -> C ABI) against
the committed outputs of the unmodified reference and against the numpy oracle.

Tolerance protocol: with ref64 = the reference run in float64,
the product must be at most as far from ref64 as 2x the reference's own float32
run is (plus one float32 ulp of the dB range), shapes bit-exact, exact-value
known answers exact.
"""
import numpy as np
import pytest
import torch

import golden_cases as gc
from oracle import frontend_np as onp

pytestmark = pytest.mark.gpu

LOBE_CASES = [n for n, s in gc.CASES.items() if s["kind"] in ("fbank", "mfcc", "fbank+norm")]
OP_CASES = [n for n, s in gc.CASES.items()
            if s["kind"] in ("stft", "magnitude", "filterbank", "dct", "deltas", "context")]
SEQ_CASES = [n for n, s in gc.CASES.items() if s["kind"] == "norm_seq"]
ULP = 1e-5  # ~ one float32 ulp at |dB| ~ 100


@pytest.fixture(scope="module")
def sb():
    assert torch.cuda.is_available(), "these tests need the B200"
    import speechbrain_b200

    return speechbrain_b200


def dev(x):
    return torch.from_numpy(np.ascontiguousarray(x)).cuda()


def _build_lobe(sb, spec):
    cls = sb.MFCC if spec["kind"] == "mfcc" else sb.Fbank
    return cls(**spec["kwargs"])


@pytest.mark.parametrize("name", LOBE_CASES)
def test_lobe_vs_reference(sb, golden, name):
    spec = gc.CASES[name]
    x = spec["make_input"]()
    mod = _build_lobe(sb, spec)
    # host constants are built with the same torch calls as the reference
    assert np.array_equal(mod.compute_fbanks.f_central.numpy(), golden[f"{name}/fc"])
    assert np.array_equal(mod.compute_fbanks.band.numpy(), golden[f"{name}/band"])
    assert np.array_equal(mod.compute_STFT.window.numpy(), golden[f"{name}/window"])
    got_t = mod(dev(x))
    got = got_t.cpu().numpy()
    ref32 = golden[f"{name}/ref32"]
    assert got.shape == ref32.shape and got.dtype == np.float32
    assert np.isfinite(got).all()
    mult = 30.0 if (spec["kind"] == "mfcc") else 1.0  # DCT sums ~23 dB values
    if f"{name}/ref64" in golden.files:
        ref64 = golden[f"{name}/ref64"]
        noise = np.abs(ref32 - ref64).max()
        err = np.abs(got - ref64).max()
        assert err <= 2 * noise + mult * ULP, (err, noise)
        assert np.abs(got - ref64).mean() <= 2 * np.abs(ref32 - ref64).mean() + ULP
    assert np.abs(got - ref32).max() < 2e-3 * mult
    if name == "cfg1_fbank40_8x1s":  # north_star gate on the CPU-runnable config
        assert np.abs(got - ref32).max() <= 1e-4
    if spec["kind"] == "fbank+norm":
        norm = sb.InputNormalization(**spec["norm"])
        lens = torch.tensor(spec["lengths"], dtype=torch.float32).cuda()
        out = norm(got_t, lens)
        if spec["norm"]["norm_type"] == "sentence":
            assert out.data_ptr() == got_t.data_ptr()  # in place, like the reference
        out = out.cpu().numpy()
        # end to end against the reference: feature error (dB) is divided by the
        # per-channel std, so the bound is the dB bound over the smallest std
        n32 = golden[f"{name}/norm32"]
        assert np.abs(out - n32).max() < 2e-3
        assert np.abs(out - n32).mean() < 2e-5
        # and against the oracle normalising OUR features (isolates the norm kernels)
        o = onp.InputNormalizationOracle(**spec["norm"])(got.astype(np.float64), spec["lengths"])
        assert np.abs(out - o).max() < 2e-5


@pytest.mark.parametrize("name", OP_CASES)
def test_op_vs_reference(sb, golden, name):
    spec = gc.CASES[name]
    x = spec["make_input"]()
    ref = golden[f"{name}/ref32"]
    kind, kw = spec["kind"], dict(spec["kwargs"])
    if kind == "stft":
        got = sb.STFT(**kw)(dev(x))
        tol = 2e-4 * np.abs(ref).max()
    elif kind == "magnitude":
        got = sb.spectral_magnitude(dev(x), **kw)
        tol = 2e-6 * max(1.0, np.abs(ref).max())
    elif kind == "filterbank":
        got = sb.Filterbank(**kw)(dev(x))
        tol = 1e-4 if kw.get("log_mel", True) else 1e-5 * np.abs(ref).max()
    elif kind == "dct":
        got = sb.DCT(**kw)(dev(x))
        tol = 1e-5 * np.abs(ref).max()
    elif kind == "deltas":
        got = sb.Deltas(**kw)(dev(x))
        tol = 1e-5
    else:
        got = sb.ContextWindow(**kw)(dev(x))
        tol = 0.0
    got = got.cpu().numpy()
    assert got.shape == ref.shape
    assert np.abs(got - ref).max() <= tol, np.abs(got - ref).max()


@pytest.mark.parametrize("name", SEQ_CASES)
def test_norm_sequence_vs_reference(sb, golden, name):
    spec = gc.CASES[name]
    norm = sb.InputNormalization(**spec["norm"])
    for i, st in enumerate(spec["steps"]):
        x = gc.synth_feats(st["seed"], spec["shape"])
        norm.train(st["training"])
        kw = {"epoch": st["epoch"]}
        if st["spk_ids"] is not None:
            kw["spk_ids"] = torch.tensor(st["spk_ids"]).unsqueeze(1)
        xin = dev(x)
        y = norm(xin, torch.tensor(st["lengths"], dtype=torch.float32).cuda(), **kw)
        inplace = spec["norm"]["norm_type"] in ("sentence", "speaker")
        assert (y.data_ptr() == xin.data_ptr()) == inplace
        if not inplace:
            assert np.array_equal(xin.cpu().numpy(), x)  # input left intact
        ref = golden[f"{name}/step{i}"]
        got = y.cpu().numpy()
        assert got.shape == ref.shape
        assert np.abs(got - ref).max() < 2e-5 * max(1.0, np.abs(ref).max()), (i, np.abs(got - ref).max())
        if spec["norm"]["norm_type"] == "global":
            assert norm.count == int(golden[f"{name}/step{i}_count"])
            gm = golden[f"{name}/step{i}_glob_mean"]
            np.testing.assert_allclose(np.broadcast_to(norm.glob_mean.cpu().numpy(), gm.shape), gm,
                                       rtol=1e-5, atol=1e-5)


def test_speaker_mode_state_round_trip(sb, tmp_path):
    """Speaker statistics live in device tables behind spk_dict_mean / spk_dict_std: the
    reference's attributes, checkpoint file and .to() keep working, table growth included."""
    g = torch.Generator().manual_seed(3)
    a = sb.InputNormalization(norm_type="speaker")
    lens = torch.tensor([1.0, 0.7, 0.9, 0.5]).cuda()
    for step in range(12):   # 4 new speakers per step: the 16-row table has to grow
        x = torch.randn(4, 50, 24, generator=g).cuda()
        spk = torch.arange(4 * step, 4 * step + 4).unsqueeze(1)
        a(x, lens, spk_ids=spk)
    assert len(a.spk_dict_mean) == 48 and a.spk_dict_count[5] == 1
    x = torch.randn(4, 50, 24, generator=g).cuda()
    spk = torch.tensor([[5], [5], [40], [0]])
    ya = a(x.clone(), lens, spk_ids=spk)
    assert a.spk_dict_count[5] == 3 and a.weight == 0.5     # the last update was speaker 0's second
    # the dictionary values are the live table rows
    assert a.spk_dict_mean[5].data_ptr() == a._spk_tab[0][a._spk_row[5]].data_ptr()
    path = str(tmp_path / "normalizer.ckpt")
    a._save(path)
    state = torch.load(path)
    assert set(state) == {"count", "glob_mean", "glob_std", "spk_dict_mean", "spk_dict_std",
                          "spk_dict_count"} and len(state["spk_dict_mean"]) == 48
    b = sb.InputNormalization(norm_type="speaker")
    b._load(path, device="cuda")
    assert b.spk_dict_count == a.spk_dict_count
    for k in a.spk_dict_mean:
        assert torch.equal(a.spk_dict_mean[k], b.spk_dict_mean[k].cuda())
    x2 = torch.randn(4, 50, 24, generator=g).cuda()
    spk2 = torch.tensor([[7], [100], [7], [5]])
    a.to("cpu").to("cuda")          # statistics follow the module (:1194-1203)
    assert torch.equal(a(x2.clone(), lens, spk_ids=spk2), b(x2.clone(), lens, spk_ids=spk2))
    a.eval(), b.eval()
    spk3 = torch.tensor([[7], [999], [100], [5]])            # 999 was never seen: own statistics
    ya, yb = a(x2.clone(), lens, spk_ids=spk3), b(x2.clone(), lens, spk_ids=spk3)
    assert torch.equal(ya, yb) and 999 not in a.spk_dict_mean
    own = sb.InputNormalization(norm_type="sentence")(x2.clone(), lens)
    assert torch.equal(ya[1], own[1])


# ---- known answers of reference tests/unittests/test_features.py, on the GPU ----
def test_ka_deltas_context(sb):
    ones = torch.ones(10, 101, 20, device="cuda")
    assert torch.equal(sb.Deltas(input_size=20)(ones), torch.zeros_like(ones))
    x = torch.tensor([1.0, 2.0, 3.0], device="cuda").view(1, 3, 1)
    want = torch.tensor([[0, 1, 2], [1, 2, 3], [2, 3, 0]], device="cuda").float().unsqueeze(0)
    assert torch.equal(sb.ContextWindow(left_frames=1, right_frames=1)(x), want)
    r = torch.rand(2, 10, 5, device="cuda")
    assert torch.equal(sb.ContextWindow()(r), r)


def test_ka_filterbank(sb):
    fb = sb.Filterbank()
    out = fb(torch.zeros(10, 101, 201, device="cuda"))
    assert torch.equal(out, torch.full_like(out, -100.0))
    assert torch.equal(fb._amplitude_to_DB(torch.zeros(1, 1, 1, device="cuda")),
                       torch.tensor([[[-100.0]]], device="cuda"))
    a = torch.rand(1, 101, 201, device="cuda") * 10
    b = torch.rand(1, 101, 201, device="cuda")
    both = fb(torch.cat([a, b]))
    assert torch.sum(torch.abs(fb(a)[0] - both[0])) < 8e-5
    assert torch.sum(torch.abs(fb(b)[0] - both[1])) < 8e-5
    x = torch.rand(10, 101, 201, device="cuda")
    two = fb(torch.stack((x, x), -1))[..., 0]
    assert torch.sum(fb(x) - two) < 5e-5


def test_ka_input_normalization(sb):
    x = torch.tensor([1.0, 2, 3, 0, 0, 0], device="cuda").view(1, 6, 1)
    out = sb.InputNormalization()(x, torch.tensor([0.5], device="cuda")).squeeze()
    assert torch.equal(out, torch.tensor([-1.0, 0, 1, -2, -2, -2], device="cuda"))


def test_shapes_and_errors(sb):
    x = torch.randn(10, 16000, device="cuda")
    assert sb.STFT(sample_rate=16000)(x).shape == (10, 101, 201, 2)
    assert sb.Fbank()(x).shape == (10, 101, 40)
    assert sb.MFCC()(x).shape == (10, 101, 660)
    assert sb.Fbank()(torch.randn(2, 0, device="cuda")).shape == (2, 1, 40)
    with pytest.raises(RuntimeError):
        sb.Fbank()(torch.randn(2, 1600))  # CPU tensor: no fallback
    with pytest.raises(ValueError):
        sb.DCT(input_size=10, n_out=20)
    half = sb.Fbank()(x.half())
    assert half.dtype == torch.float32


@pytest.mark.parametrize("B,L,kw", [(32, 256000, {"n_mels": 80}), (3, 48000 + 77, {"n_mels": 80}),
                                    (2, 480000, {"n_mels": 80})])
def test_full_size_properties(sb, B, L, kw):
    """BASELINE-size checks through size-independent properties: batch
    independence, time-shift equivariance by whole hops, clamp floor."""
    g = torch.Generator(device="cuda").manual_seed(5)
    wav = torch.randn(B, L, device="cuda", generator=g) * 0.1
    mod = sb.Fbank(**kw)
    full = mod(wav)
    assert full.shape == (B, 1 + L // 160, 80) and torch.isfinite(full).all()
    one = mod(wav[1:2].clone())
    assert torch.equal(one[0], full[1])  # utterances are independent, bit for bit
    # frames away from the edges depend only on their own 400 samples
    shifted = mod(wav[:1, 160 * 7:].clone())
    a, b_ = full[0, 7 + 3:-3], shifted[0, 3:-3]
    floor = max(full[0].max(), shifted[0].max()) - 80 + 1e-3
    keep = (a > floor) & (b_ > floor)
    assert torch.equal(a[keep], b_[keep])
    assert (full.amax(dim=(1, 2)) - full.amin(dim=(1, 2)) <= 80.0 + 1e-4).all()


def test_full_size_properties_cfg3_cfg4_cfg5(sb):
    """The other BASELINE.json configs at their full sizes, through properties that do not need
    the oracle: batch independence (bit for bit), linearity of the statistics, in-place
    contract, zero mean / unit variance over the valid frames."""
    g = torch.Generator(device="cuda").manual_seed(7)
    # cfg3: MFCC (23 mel -> 20 DCT -> deltas -> context 5+5), 64 x 4 s
    wav = torch.randn(64, 64000, device="cuda", generator=g) * 0.1
    mf = sb.MFCC()
    full = mf(wav)
    assert full.shape == (64, 401, 660) and torch.isfinite(full).all()
    assert torch.equal(mf(wav[5:6].clone())[0], full[5])
    ctx = full.view(64, 401, 60, 11)                       # [.., feature, context]: f * C + c
    assert torch.equal(ctx[:, 5:-5, :, 5 + 3], ctx[:, 8:-2, :, 5])   # window slot c = frame t + c - 5
    assert bool((ctx[:, 0, :, :5] == 0).all()) and bool((ctx[:, -1, :, 6:] == 0).all())   # zero padding
    # cfg4: Fbank-80 + sentence mean normalisation, 256 x 3 s (pair() is a no-op shim now)
    wav = torch.randn(256, 48000, device="cuda", generator=g) * 0.1
    lens = torch.linspace(0.5, 1.0, 256, device="cuda")
    for i in range(256):
        wav[i, int(round(float(lens[i]) * 48000)):] = 0
    fb, norm = sb.Fbank(n_mels=80), sb.InputNormalization(norm_type="sentence", std_norm=False)
    feats = fb(wav)
    out = norm(feats, lens)
    assert out.data_ptr() == feats.data_ptr() and out.shape == (256, 301, 80)
    n_valid = torch.round(lens * 301).long()
    for i in (0, 100, 255):
        assert float(out[i, :n_valid[i]].mean(0).abs().max()) < 2e-5
    fb2, norm2 = sb.Fbank(n_mels=80), sb.InputNormalization(norm_type="sentence", std_norm=False)
    sb.pair(fb2, norm2)
    assert torch.equal(norm2(fb2(wav), lens), out)


def test_finished_output_and_tile_records_match_generic_kernels(sb):
    """Tile-record mode of the fused kernel (clamp + per-tile moments inside the producing launch) ==
    the un-clamped kernel followed by the generic clamp / statistics kernels, with silence tails
    (constant tiles at the top_db floor), an all-zero utterance, a quiet stretch INSIDE an
    utterance (non-constant tiles below the floor) and valid lengths that cut tiles."""
    from speechbrain_b200 import ops

    rng = np.random.default_rng(5)
    wav = (rng.standard_normal((6, 3 * 16000 + 77)) * 0.1).astype(np.float32)
    wav[1, 30000:] = 0
    wav[2, 100:] = 0
    wav[3, :] = 0
    wav[4, 20000:20800] *= 1e-5
    wav[5, 5000:9000] *= 1e-6
    x = dev(wav)
    fb = sb.Fbank(n_mels=80)
    feats = fb(x)
    ts = getattr(feats, "_sbfe_tile_stats", None)
    assert ts is not None and ts.frames == 32 and ts.version == feats._version
    raw, utt_max = ops.fbank_forward(fb._frontend_tables(), x)
    ref = ops.post_forward(raw, utt_max, 80.0)
    assert torch.equal(feats, ref)
    assert float((ref - raw).abs().max()) > 1.0          # the clamp did something
    for lens in ([1.0, 0.73, 0.5, 0.051, 1.0, 0.9], [0.999, 0.32, 0.0, 1.0, 0.5, 0.0107],
                 [32 / 301, 64 / 301, 31 / 301, 33 / 301, 1 / 301, 300 / 301]):
        lens = torch.tensor(lens).cuda()
        m1, s1 = ops.norm_stats_tiles(feats, lens, ts)
        m2, s2 = ops.norm_stats(feats, lens)
        assert torch.allclose(m1, m2, rtol=1e-5, atol=2e-5, equal_nan=True), float((m1 - m2).abs().max())
        assert torch.allclose(s1, s2, rtol=2e-5, atol=2e-5, equal_nan=True), float((s1 - s2).abs().max())
    # the hand-over is one-shot and version-checked
    lens = torch.tensor([1.0, 0.73, 0.5, 0.051, 1.0, 0.9]).cuda()
    m2, s2 = ops.norm_stats(feats, lens)
    norm = sb.InputNormalization(norm_type="sentence")
    want = (feats - m2[:, None, :]) / s2[:, None, :]
    got = norm(feats, lens)
    assert got.data_ptr() == feats.data_ptr()
    assert getattr(got, "_sbfe_tile_stats", None) is None
    assert torch.allclose(got, want, rtol=1e-4, atol=1e-4)
    # the other modes take their statistics from the records as well
    for mode in ("batch", "global", "speaker"):
        f1, f2 = fb(x), fb(x)
        f2._sbfe_tile_stats = None                        # generic statistics kernels
        kw = {"spk_ids": torch.tensor([3, 5, 3, 7, 5, 3])} if mode == "speaker" else {}
        a = sb.InputNormalization(norm_type=mode)(f1, lens, **kw)
        b_ = sb.InputNormalization(norm_type=mode)(f2, lens, **kw)
        torch.testing.assert_close(a, b_, rtol=1e-4, atol=1e-4)


def test_tile_record_mode_is_deterministic_and_its_records_are_exact(sb):
    """Mode (b) of sbfe_fbank_forward straight through the C ABI (producing kernel + the clamp
    launch that reads the records) on batches with more tiles than one wave of CTAs, many clamped
    and constant tiles, tile groups that end inside an utterance: same bits every time,
    == un-clamped kernel + generic clamp; the records hold the moments of the un-clamped tiles
    (consumers check each column's minimum against the floor, norm.cu)."""
    import ctypes as C

    from speechbrain_b200 import _lib, ops

    rng = np.random.default_rng(17)
    for B, L in ((48, 16 * 16000), (3, 16000 + 53), (1, 160 * 31), (300, 4000)):
        wav = (rng.standard_normal((B, L)) * 0.1).astype(np.float32)
        n = rng.integers(min(800, L // 2), L, B)
        for i in range(B):
            wav[i, n[i]:] = 0                               # silence tails: constant tiles
            wav[i, n[i] // 2: n[i] // 2 + L // 64] *= 1e-6    # quiet stretch: tiles below the floor
        x = dev(wav)
        fb = sb.Fbank(n_mels=80)
        tab = fb._frontend_tables()
        raw, utt_max = ops.fbank_forward(tab, x)
        ref = ops.post_forward(raw, utt_max, 80.0)
        desc, _ = tab.on(x.device)
        L_ = _lib.lib()
        T = 1 + L // 160
        tiles = (T + 31) // 32
        words = int(L_.sbfe_fbank_tile_stats_floats(C.byref(desc)))
        first = None
        for trial in range(6):
            out = torch.empty(B, T, 80, device="cuda")
            rec = torch.full((B, tiles, words), float("nan"), device="cuda")
            rc = L_.sbfe_fbank_forward(C.byref(desc), C.c_void_p(x.data_ptr()), B, L, L,
                                       C.c_void_p(out.data_ptr()), None, C.c_void_p(rec.data_ptr()),
                                       80.0, None, 0,
                                       C.c_void_p(torch.cuda.current_stream().cuda_stream))
            assert rc == 0
            assert torch.equal(out, ref), (B, L, trial)
            assert bool(torch.isfinite(rec[..., :160]).all())
            if first is None:
                first = rec.clone()
            assert torch.equal(rec[..., :160], first[..., :160])      # per-tile mean / M2
        # the records' moments (of the UN-clamped values) against a direct computation in float64
        r64 = raw.double()
        for j in sorted({0, tiles // 2, tiles - 1}):
            seg = r64[:, 32 * j: min(32 * (j + 1), T)]
            mu = seg.mean(1)
            m2 = ((seg - mu[:, None]) ** 2).sum(1)
            assert torch.allclose(first[:, j, :80].double(), mu, rtol=1e-5, atol=1e-4)
            assert torch.allclose(first[:, j, 80:160].double(), m2, rtol=1e-4, atol=1e-2)


def test_in_place_ops_void_the_tile_records(sb):
    """an in-place op of this package between Fbank and InputNormalization (here
    SpecAugment without a warp: masks written through ctypes) must void the hand-over - the
    statistics then come from the generic kernels and see the masks."""
    rng = np.random.default_rng(33)
    wav = dev((rng.standard_normal((4, 32000)) * 0.1).astype(np.float32))
    lens = torch.tensor([1.0, 0.8, 0.6, 0.9]).cuda()
    fb = sb.Fbank(n_mels=80)
    aug = sb.SpecAugment(time_warp=False, freq_mask=True, time_mask=True, freq_mask_width=(10, 20),
                         time_mask_width=(20, 40), n_freq_mask=2, n_time_mask=2)
    aug.train()
    torch.manual_seed(3)
    f = fb(wav)
    assert f._sbfe_tile_stats is not None
    g = aug(f)
    assert g.data_ptr() == f.data_ptr() and float((g == 0).float().mean()) > 0.05
    masked = g.clone()
    got = sb.InputNormalization(norm_type="sentence")(g, lens)
    want = sb.InputNormalization(norm_type="sentence")(masked, lens)     # no records: generic path
    assert torch.equal(got, want)
    # plain torch in-place edits are caught by the same version check
    f = fb(wav)
    f += 1.0
    keep = f.clone()
    assert torch.equal(sb.InputNormalization(norm_type="sentence")(f, lens),
                       sb.InputNormalization(norm_type="sentence")(keep, lens))
    assert sb.pair(fb, sb.InputNormalization()) is fb                     # round-1 shim: a no-op


def test_ragged_ingest_is_bit_identical(sb):
    """RaggedBatch (packed valid samples, zero-fill in the kernel) == the padded batch."""
    rng = np.random.default_rng(9)
    sizes = [48000, 31999, 160, 47777, 1, 20480]
    wavs = [torch.from_numpy((rng.standard_normal(n) * 0.1).astype(np.float32)) for n in sizes]
    rb = sb.RaggedBatch.from_list(wavs, pin_memory=False).to("cuda")
    padded = rb.to_padded()
    assert padded.shape == (6, 48000)
    fb = sb.Fbank(n_mels=80)
    a, b = fb(rb), fb(padded)
    assert a.shape == b.shape == (6, 301, 80) and torch.equal(a, b)
    mf = sb.MFCC()
    assert torch.equal(mf(rb), mf(padded))
    norm = sb.InputNormalization(norm_type="sentence")
    # (the 1-sample utterance has 0 valid frames: NaN statistics, like the reference)
    assert torch.allclose(norm(fb(rb), rb.lengths), norm(fb(padded), rb.lengths), rtol=0, atol=0,
                          equal_nan=True)
    # from_padded round trip
    host = padded.cpu()
    rb2 = sb.RaggedBatch.from_padded(host, rb.lengths.cpu(), pin_memory=False).to("cuda")
    assert torch.equal(fb(rb2), b)


def test_all_zero_tiles_take_the_constant_path(sb):
    """Padding behind an utterance / digital silence: the fused kernel writes the floor without
    transforming (fbank400_v3.cu); results must equal what the transform of zeros gives."""
    from speechbrain_b200 import ops

    rng = np.random.default_rng(21)
    wav = (rng.standard_normal((5, 48000)) * 0.1).astype(np.float32)
    wav[1, 10000:] = 0                       # padded tail
    wav[2] = 0                               # silent utterance
    wav[3] = 0
    wav[3, 30007] = 1e-3                     # one sample deep inside an otherwise silent tile
    wav[4, 12000:30000] = -0.0               # silence in the middle, negative zeros
    fb = sb.Fbank(n_mels=80)
    feats, keys = ops.fbank_forward(fb._frontend_tables(), dev(wav))    # un-clamped
    got = feats.cpu().numpy()
    consts = (fb.compute_fbanks.f_central.numpy(), fb.compute_fbanks.band.numpy())
    want = onp.fbank(wav, n_mels=80, dtype=np.float64, mel_params=consts,
                     window=fb.compute_STFT.window.numpy())          # with the top_db clamp
    assert np.abs(fb(dev(wav)).cpu().numpy() - want).max() < 1e-3
    t = np.arange(301)
    silent = lambda lo, hi: (t * 160 - 200 >= lo) & (t * 160 + 200 <= hi)   # frames inside [lo, hi)
    assert (got[1][silent(10000, 48000)] == -100.0).all()
    assert (got[2] == -100.0).all()
    touched = (t * 160 - 200 <= 30007) & (30007 < t * 160 + 200)    # frames whose window holds it
    assert touched.sum() in (2, 3) and (got[3][~touched] == -100.0).all()
    assert (got[3][touched].max(axis=1) > -100.0).all()
    assert (got[4][silent(12000, 30000)] == -100.0).all()
    assert ops.decode_utt_max(keys)[2].item() == -100.0
    # the statistics hand-over sees constant tiles as (mean = floor, M2 = 0)
    lens = torch.tensor([1.0, 0.5, 1.0, 1.0, 1.0]).cuda()
    a = sb.InputNormalization(norm_type="sentence", std_norm=False)(fb(dev(wav)), lens)
    f2 = fb(dev(wav))
    f2._sbfe_tile_stats = None
    b = sb.InputNormalization(norm_type="sentence", std_norm=False)(f2, lens)
    torch.testing.assert_close(a, b, rtol=0, atol=2e-4)
    # linear mel: zeros stay zeros
    lin = sb.Fbank(n_mels=40)
    lin.compute_fbanks.log_mel = False
    lin._tables = None
    out = lin(dev(wav))
    assert bool((out[2] == 0).all())


@pytest.mark.parametrize("n_mels", [96, 128, 161])
def test_more_than_eighty_filters(sb, n_mels):
    """n_mels > 80: the fused kernel's mel stage takes several passes of 80 filters (the
    tile-invariant plan only covers the first); checked against the float64 oracle."""
    rng = np.random.default_rng(n_mels)
    wav = (rng.standard_normal((3, 16000 + 53)) * 0.1).astype(np.float32)
    wav[1, 9000:] = 0
    fb = sb.Fbank(n_mels=n_mels)
    got = fb(dev(wav)).cpu().numpy()
    consts = (fb.compute_fbanks.f_central.numpy(), fb.compute_fbanks.band.numpy())
    want = onp.fbank(wav, n_mels=n_mels, dtype=np.float64, mel_params=consts,
                     window=fb.compute_STFT.window.numpy())
    assert got.shape == want.shape == (3, 101, n_mels)
    assert np.abs(got - want).max() < 1e-3   # dB; fp32 FFT + lg2.approx vs float64
    norm = sb.InputNormalization(norm_type="sentence")
    lens = torch.tensor([1.0, 0.6, 1.0]).cuda()
    a = norm(fb(dev(wav)), lens).cpu().numpy()
    b = onp.InputNormalizationOracle(norm_type="sentence")(want.copy(), [1.0, 0.6, 1.0])
    # a column the top_db clamp made constant has zero variance: (rounding noise) / eps is
    # arbitrary in every implementation, the reference's float32 included - not compared
    live = want[:, :60].std(axis=1) > 1e-3                    # [B, M]
    assert live.mean() > 0.9
    assert np.abs(a - b)[np.broadcast_to(live[:, None, :], a.shape)].max() < 1e-3


def test_pcm16_ingest_is_bit_identical(sb, golden):
    """16-bit PCM kept as int16 until the fused kernel stages it (scale 2^-15 folded into the
    window) == the same samples converted to float32 on the host, bit for bit."""
    from speechbrain_b200 import ops

    rng = np.random.default_rng(10)
    sizes = [48000, 31999, 160, 47777, 1, 0, 20481]
    pcm = [torch.from_numpy(np.clip(np.rint(rng.standard_normal(n) * 3000), -32768, 32767).astype(np.int16))
           for n in sizes]
    pcm[0][:5] = torch.tensor([32767, -32768, 0, 1, -1], dtype=torch.int16)
    rb16 = sb.RaggedBatch.from_list(pcm, pin_memory=False)
    assert rb16.packed.dtype == torch.int16 and rb16.nbytes < 2 * sum(sizes) + 200
    rb32 = sb.RaggedBatch.from_list([w.float() / 32768 for w in pcm], pin_memory=False)
    rb16, rb32 = rb16.to("cuda"), rb32.to("cuda")
    assert torch.equal(rb16.to_padded(), rb32.to_padded())
    for mod in (sb.Fbank(n_mels=80), sb.Fbank(n_mels=40, deltas=True), sb.MFCC()):
        assert torch.equal(mod(rb16), mod(rb32)), type(mod).__name__
    fb, norm = sb.Fbank(n_mels=80), sb.InputNormalization(norm_type="sentence")
    assert torch.allclose(norm(fb(rb16), rb16.lengths), norm(fb(rb32), rb32.lengths), rtol=0, atol=0,
                          equal_nan=True)
    # a buffer that starts at an odd sample (2-byte aligned only): the un-vectorised staging path
    shifted = torch.zeros(rb16.packed.numel() + 1, dtype=torch.int16, device="cuda")
    shifted[1:] = rb16.packed
    odd = sb.RaggedBatch(shifted[1:], rb16.offsets, rb16.nsamples, rb16.max_len, rb16.lengths)
    assert odd.packed.data_ptr() % 4 == 2
    assert torch.equal(fb(odd), fb(rb32))
    # real speech from the reference's test sample, as the int16 PCM the file holds
    import golden_cases as gcases

    speech = torch.from_numpy(np.load(gcases.SPEECH_NPZ)["pcm"].astype(np.int16))
    a = fb(sb.RaggedBatch.from_list([speech, speech[:9000]], pin_memory=False).to("cuda"))
    b = fb(sb.RaggedBatch.from_list([speech.float() / 32768, speech[:9000].float() / 32768],
                                    pin_memory=False).to("cuda"))
    assert torch.equal(a, b)
    # a scale that is not a power of two still works (no longer bit-identical, 1-ulp level)
    rb16.scale = 1.0 / 30000
    want = fb(rb16.to_padded())
    torch.testing.assert_close(fb(rb16), want, rtol=0, atol=2e-4)


# ---- torch.jit.trace, as the reference's unit tests do (tests/unittests/test_features.py:
# 14, 39, 56, 66, 97, 107): the traced module must replay the CUDA kernels, not a constant ----
def test_modules_survive_jit_trace(sb):
    g = torch.Generator().manual_seed(11)
    wav = torch.randn(4, 16000, generator=g).cuda()
    wav2 = torch.randn(4, 16000, generator=g).cuda()
    spec = torch.rand(4, 101, 201, generator=g).cuda()
    spec2 = torch.rand(4, 101, 201, generator=g).cuda()
    feat = torch.rand(4, 101, 40, generator=g).cuda()
    feat2 = torch.rand(4, 101, 40, generator=g).cuda()
    cases = [
        (sb.STFT(sample_rate=16000), wav, wav2),
        (sb.Filterbank(), spec, spec2),
        (sb.DCT(input_size=40), feat, feat2),
        (sb.Deltas(input_size=40), feat, feat2),
        (sb.ContextWindow(left_frames=2, right_frames=3), feat, feat2),
        (sb.Fbank(), wav, wav2),
        (sb.MFCC(), wav, wav2),
    ]
    for mod, a, b in cases:
        traced = torch.jit.trace(mod, (a,))
        assert "sbfe::module_call" in str(traced.graph), type(mod).__name__
        # a different input through the traced graph == the eager module on that input
        assert torch.equal(traced(b), mod(b)), type(mod).__name__

    lens = torch.tensor([1.0, 0.8, 0.5, 0.3]).cuda()
    for kind in ("sentence", "batch"):
        norm = sb.InputNormalization(norm_type=kind)
        traced = torch.jit.trace(norm, (feat.clone(), lens))
        x = feat2.clone()
        y = traced(x, lens)
        want = sb.InputNormalization(norm_type=kind)(feat2.clone(), lens)
        assert torch.equal(y, want), kind
        if kind == "sentence":  # the in-place contract survives tracing
            assert torch.equal(x, want)
        else:
            assert torch.equal(x, feat2)


# ---- guard bands: the kernels write exactly their outputs (compute-sanitizer is not available
# on the GPU pool, so out-of-bounds stores are caught with canaries around every output) ----
@pytest.mark.parametrize("B,L,M", [(3, 16000 + 53, 80), (2, 5119, 23), (1, 160 * 33, 40), (5, 0, 80),
                                   (2, 48000, 128)])
def test_outputs_stay_inside_their_buffers(sb, B, L, M):
    import ctypes as C

    from speechbrain_b200 import _lib, ops

    G, CAN = 4096, 1234567.0
    L_ = _lib.lib()
    fb = sb.Fbank(n_mels=M)
    tab = fb._frontend_tables()
    desc, _ = tab.on(torch.device("cuda", 0))
    T = 1 + L // 160
    wav = torch.randn(B, max(L, 1), device="cuda")[:, :L].contiguous() * 0.1
    frames = int(L_.sbfe_fbank_tile_frames(C.byref(desc)))
    words = int(L_.sbfe_fbank_tile_stats_floats(C.byref(desc)))
    tiles = (T + frames - 1) // frames
    n_out, n_ts = B * T * M, B * tiles * words
    buf_o = torch.full((n_out + 2 * G,), CAN, device="cuda")
    buf_t = torch.full((n_ts + 2 * G,), CAN, device="cuda")
    s = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = L_.sbfe_fbank_forward(C.byref(desc), C.c_void_p(wav.data_ptr()), B, L, max(L, 1),
                               C.c_void_p(buf_o[G:].data_ptr()), None, C.c_void_p(buf_t[G:].data_ptr()),
                               80.0, None, 0, s)
    assert rc == 0
    torch.cuda.synchronize()
    for buf, n in ((buf_o, n_out), (buf_t, n_ts)):
        assert bool((buf[:G] == CAN).all()) and bool((buf[G + n:] == CAN).all())
    out = buf_o[G:G + n_out].view(B, T, M)
    assert not bool((out == CAN).any())
    raw, keys = ops.fbank_forward(tab, wav)
    assert torch.equal(out, ops.post_forward(raw, keys, 80.0))
    # normalisation in place from the tile records (beside the clamp launch of a second call,
    # and the ordinary way: same bits), SpecAugment out of place
    ts = ops.TileStats(buf_t[G:G + n_ts].view(B, tiles, words), frames, 80.0)
    lens = torch.linspace(0.5, 1.0, B, device="cuda")
    plain = ops.norm_sentence_tiles(out.clone(), lens, ts)
    rc = L_.sbfe_fbank_forward(C.byref(desc), C.c_void_p(wav.data_ptr()), B, L, max(L, 1),
                               C.c_void_p(buf_o[G:].data_ptr()), None, C.c_void_p(buf_t[G:].data_ptr()),
                               80.0, None, 0, s)
    assert rc == 0
    ops.norm_sentence_tiles(out, lens, ts, out=out, beside_clamp=True)
    torch.cuda.synchronize()
    assert torch.equal(torch.nan_to_num(out, nan=7.0), torch.nan_to_num(plain, nan=7.0))  # T = 1: NaN std
    torch.cuda.synchronize()
    assert bool((buf_o[:G] == CAN).all()) and bool((buf_o[G + n_out:] == CAN).all())
    if T > 12:
        n_aug = B * T * M
        buf_a = torch.full((n_aug + 2 * G,), CAN, device="cuda")
        z = torch.zeros(B, 1, dtype=torch.int64, device="cuda")
        rc = L_.sbfe_specaugment_forward(C.c_void_p(out.data_ptr()), B, T, M, T // 2, T // 2 + 1, 0,
                                         C.c_void_p(z.data_ptr()), C.c_void_p((z + 3).data_ptr()), 1,
                                         C.c_void_p(z.data_ptr()), C.c_void_p((z + 2).data_ptr()), 1,
                                         None, 0.0, C.c_void_p(buf_a[G:].data_ptr()), s)
        assert rc == 0
        torch.cuda.synchronize()
        assert bool((buf_a[:G] == CAN).all()) and bool((buf_a[G + n_aug:] == CAN).all())
        assert not bool((buf_a[G:G + n_aug] == CAN).any())


def test_post_output_stays_inside_its_buffer(sb):
    import ctypes as C

    from speechbrain_b200 import _lib

    G, CAN = 4096, 1234567.0
    L_ = _lib.lib()
    for B, T, F, n_dct, deltas, left, right in [(3, 101, 23, 20, 1, 5, 5), (2, 37, 40, 0, 1, 0, 0),
                                                (1, 7, 23, 20, 0, 2, 3), (4, 64, 40, 0, 0, 1, 1)]:
        x = torch.randn(B, T, F, device="cuda")
        F0 = n_dct if n_dct else F
        F_out = F0 * (3 if deltas else 1) * (left + right + 1)
        dct = sb.DCT(input_size=F, n_out=n_dct).matrix_on(x.device) if n_dct else None
        buf = torch.full((B * T * F_out + 2 * G,), CAN, device="cuda")
        rc = L_.sbfe_post_forward(C.c_void_p(x.data_ptr()), B, T, F, None, 80.0,
                                  C.c_void_p(dct.data_ptr()) if n_dct else None, n_dct, deltas, 5,
                                  int(bool(left or right)), left, right, C.c_void_p(buf[G:].data_ptr()),
                                  C.c_void_p(torch.cuda.current_stream().cuda_stream))
        assert rc == 0
        torch.cuda.synchronize()
        assert bool((buf[:G] == CAN).all()) and bool((buf[G + B * T * F_out:] == CAN).all())
        assert not bool((buf[G:G + B * T * F_out] == CAN).any())
