"""Round-2 GPU parity: the product against the round-2 reference fixtures (jitter, other hops,
n_fft 512, gaussian filters, normalized STFT, reference-written checkpoints) and - at the FULL
sizes of BASELINE.json's configs - directly against the numpy oracle in float64 with the
SURVEY.md 8c protocol (the product may be at most twice as far from the float64 anchor as the
float32 restatement of the reference is)."""
import numpy as np
import pytest
import torch

import golden_cases as gc
from oracle import frontend_np as onp

pytestmark = pytest.mark.gpu
ULP = 1e-5
LOBES_R2 = [n for n, s in gc.CASES_R2.items() if s["kind"] in ("fbank", "mfcc")]


@pytest.fixture(scope="module")
def sb():
    assert torch.cuda.is_available(), "these tests need the B200"
    import speechbrain_b200

    return speechbrain_b200


@pytest.fixture(scope="module")
def golden2():
    return np.load(gc.GOLDEN_R2_NPZ)


def dev(x):
    return torch.from_numpy(np.ascontiguousarray(x)).cuda()


@pytest.mark.parametrize("name", LOBES_R2)
def test_lobe_vs_reference_r2(sb, golden2, name):
    spec = gc.CASES_R2[name]
    x = spec["make_input"]()
    mod = (sb.MFCC if spec["kind"] == "mfcc" else sb.Fbank)(**spec["kwargs"])
    assert np.array_equal(mod.compute_fbanks.f_central.numpy(), golden2[f"{name}/fc"])
    assert np.array_equal(mod.compute_STFT.window.numpy(), golden2[f"{name}/window"])
    if "train_seed" in spec:      # jitter: training mode, same default-generator draw
        mod.train()
        torch.manual_seed(spec["train_seed"])
    else:
        mod.eval()
    got = mod(dev(x)).cpu().numpy()
    ref32 = golden2[f"{name}/ref32"]
    assert got.shape == ref32.shape and np.isfinite(got).all()
    mult = 30.0 if spec["kind"] == "mfcc" else 1.0
    if f"{name}/ref64" in golden2.files:
        ref64 = golden2[f"{name}/ref64"]
        noise = np.abs(ref32 - ref64).max()
        err = np.abs(got - ref64).max()
        assert err <= 2 * noise + mult * ULP, (err, noise)
        assert np.abs(got - ref64).mean() <= 2 * np.abs(ref32 - ref64).mean() + ULP
    assert np.abs(got - ref32).max() < 2e-3 * mult
    if "train_seed" in spec:      # and a second call draws again: different filters
        again = mod(dev(x)).cpu().numpy()
        assert np.abs(again - got).max() > 1e-3
        mod.eval()
        base = (sb.MFCC if spec["kind"] == "mfcc" else sb.Fbank)(
            **{k: v for k, v in spec["kwargs"].items() if k != "param_rand_factor"})
        assert torch.equal(mod(dev(x)), base(dev(x)))   # eval mode: no jitter


def test_stft_normalized_vs_reference(sb, golden2):
    for name in ("stft_normalized", "stft_normalized_nfft512"):
        spec = gc.CASES_R2[name]
        got = sb.STFT(**spec["kwargs"])(dev(spec["make_input"]())).cpu().numpy()
        ref = golden2[f"{name}/ref32"]
        assert got.shape == ref.shape
        assert np.abs(got - ref).max() <= 2e-4 * np.abs(ref).max()


def test_filterbank_gaussian_module_vs_reference(sb, golden2):
    spec = gc.CASES_R2["filterbank_gaussian_spec"]
    got = sb.Filterbank(**spec["kwargs"])(dev(spec["make_input"]())).cpu().numpy()
    ref = golden2["filterbank_gaussian_spec/ref32"]
    assert got.shape == ref.shape and np.abs(got - ref).max() < 1e-4


@pytest.mark.parametrize("norm_type,path", [("global", gc.REF_CKPT_GLOBAL), ("speaker", gc.REF_CKPT_SPEAKER)])
def test_reference_written_checkpoint_drives_the_kernels(sb, golden2, norm_type, path):
    """normalizer.ckpt written by the reference's _save, loaded by the product (as the
    Checkpointer does: _load(path, end_of_epoch, device)), then an eval step on the GPU ==
    the reference's eval step after the same state."""
    norm = sb.InputNormalization(norm_type=norm_type)
    norm._load(path, end_of_epoch=False, device="cuda")
    norm.eval()
    x = dev(gc.synth_feats(gc.CKPT_EVAL["seed"], gc.CKPT_SHAPE))
    y = norm(x, torch.tensor(gc.CKPT_EVAL["lengths"]).cuda(),
             spk_ids=torch.tensor(gc.CKPT_EVAL["spk_ids"]).unsqueeze(1))
    ref = golden2[f"ckpt_{norm_type}/eval"]
    assert np.abs(y.cpu().numpy() - ref).max() < 2e-5 * max(1.0, np.abs(ref).max())
    assert norm.count == int(golden2[f"ckpt_{norm_type}/count"])


# ---- BASELINE.json configs at FULL size against the oracle -----------------
def _protocol(got, ref64, ref32, mult=1.0):
    """SURVEY.md 8c (ii) for the small cases: no less accurate than twice the float32 restatement
    of the reference."""
    noise = np.abs(ref32 - ref64).max()
    err = np.abs(got - ref64).max()
    assert err <= 2 * noise + mult * ULP, (err, noise)
    assert np.abs(got - ref64).mean() <= 2 * np.abs(ref32 - ref64).mean() + mult * ULP
    return err, noise


def _protocol_full(got, ref64, ref32):
    """The same anchor at the BASELINE sizes (millions of values of white noise, whose spectra have
    nulls 50-60 dB below the frame's level).  There a float32 FFT's error - relative to the FRAME,
    not to the bin - shows in the deepest nulls: the reference's own ATen chain on the cfg2 batch
    is off the float64 anchor by up to 3.1e-4 dB through cuFFT and 7.0e-4 dB on the host, with
    11 / 24 of 4.1 M values beyond 1e-4; the fused kernel: 5.5e-4, 10 values; numpy's float32
    FFT is unusually tight (7.7e-5), so the small-case rule would demand more than the reference
    delivers (profiles/r2_accuracy_cfg2.txt, tools/diag_cfg2_err.py).  Stated tolerance:
    mean error <= 2 x the numpy float32 restatement's + 1e-5, at most 1e-5 of the values beyond
    1e-4 dB, none beyond 1e-3 dB."""
    err = np.abs(got - ref64)
    assert err.mean() <= 2 * np.abs(ref32 - ref64).mean() + ULP, err.mean()
    assert (err > 1e-4).mean() <= 1e-5, int((err > 1e-4).sum())
    assert err.max() <= 1e-3, err.max()


def _oracle_fbank(fb, wav, dtype, n_mels=80):
    consts = (fb.compute_fbanks.f_central.numpy(), fb.compute_fbanks.band.numpy())
    return onp.fbank(wav, n_mels=n_mels, dtype=dtype, mel_params=consts,
                     window=fb.compute_STFT.window.numpy())


def _check_sentence_norm(out, ref64, lens, std_norm=True):
    """Normalised output against the oracle normalising the float64 features; columns the
    top_db clamp made constant over the valid frames (zero variance) are not compared - there
    (rounding noise) / eps is arbitrary in every implementation, the reference included."""
    want = onp.InputNormalizationOracle(norm_type="sentence", std_norm=std_norm)(ref64.copy(), lens)
    T = ref64.shape[1]
    for b in range(ref64.shape[0]):
        n = int(np.rint(np.float32(lens[b]) * np.float32(T)))
        live = ref64[b, :n].std(axis=0) > 1e-3 if std_norm else np.ones(ref64.shape[2], bool)
        assert live.mean() > 0.9
        d = np.abs(out[b][:, live] - want[b][:, live])
        assert d.max() < 1e-3, (b, d.max())
        assert d.mean() < 2e-5


def test_cfg2_full_size_vs_oracle(sb):
    """32 x 16 s, lengths 0.6-1.0 with zero-padded tails: exactly bench.py's synth_batch."""
    import bench

    wav_t, lens_t = bench.synth_batch(0, "cpu")
    wav, lens = wav_t.numpy(), lens_t.numpy()
    fb, norm = sb.Fbank(n_mels=80), sb.InputNormalization(norm_type="sentence")
    feats = fb(wav_t.cuda())
    got = feats.cpu().numpy()
    assert got.shape == (32, 1601, 80)
    ref64 = _oracle_fbank(fb, wav, np.float64)
    ref32 = _oracle_fbank(fb, wav, np.float32)
    _protocol_full(got, ref64, ref32)
    out = norm(feats, lens_t.cuda())
    assert out.data_ptr() == feats.data_ptr()
    _check_sentence_norm(out.cpu().numpy(), ref64, lens)


def test_nfft512_full_size_vs_oracle_and_generic_path(sb):
    """Fbank(n_fft=512) on the cfg2 batch through the fused 512-point kernel (fbank512.cu: 16 x 16
    complex FFT + untangle in registers, tile records, clamp launch, one-launch normalisation)
    against the float64 oracle, and against the library's generic Stockham path on a batch slice."""
    import os

    import bench

    wav_t, lens_t = bench.synth_batch(3, "cpu")
    wav, lens = wav_t.numpy(), lens_t.numpy()
    fb = sb.Fbank(n_mels=80, n_fft=512)
    feats = fb(wav_t.cuda())
    assert getattr(feats, "_sbfe_tile_stats", None) is not None      # the fused kernel took it
    got = feats.cpu().numpy()
    assert got.shape == (32, 1601, 80)
    consts = (fb.compute_fbanks.f_central.numpy(), fb.compute_fbanks.band.numpy())
    kw = dict(n_mels=80, n_fft=512, mel_params=consts, window=fb.compute_STFT.window.numpy())
    sl = [0, 13, 31]
    ref64 = onp.fbank(wav[sl], dtype=np.float64, **kw)
    ref32 = onp.fbank(wav[sl], dtype=np.float32, **kw)
    _protocol_full(got[sl], ref64, ref32)
    # STFT(n_fft=512) is the Stockham kernel: an independent implementation inside the library
    st = sb.STFT(sample_rate=16000, n_fft=512)(wav_t[sl].cuda())
    power = (st ** 2).sum(-1)
    mel = torch.matmul(power, fb.compute_fbanks.fbank_matrix().cuda())
    x_db = 10 * torch.log10(torch.clamp(mel, min=1e-10))
    x_db = torch.maximum(x_db, x_db.amax(dim=(1, 2), keepdim=True) - 80.0)
    assert float((feats[sl] - x_db).abs().max()) < 2e-3
    out = sb.InputNormalization(norm_type="sentence")(feats, lens_t.cuda())
    full64 = onp.fbank(wav[:4], dtype=np.float64, **kw)
    _check_sentence_norm(out[:4].cpu().numpy(), full64, lens[:4])


def test_cfg3_full_size_vs_oracle(sb):
    """64 x 4 s MFCC: 23 mel -> 20 DCT -> deltas -> context 5+5 = 660."""
    wav = gc.synth_wav(703, 64, 64000)
    mf = sb.MFCC()
    got = mf(dev(wav)).cpu().numpy()
    assert got.shape == (64, 401, 660)
    consts = (mf.compute_fbanks.f_central.numpy(), mf.compute_fbanks.band.numpy())
    kw = dict(mel_params=consts, window=mf.compute_STFT.window.numpy())
    ref64 = onp.mfcc(wav, dtype=np.float64, **kw)
    ref32 = onp.mfcc(wav, dtype=np.float32, **kw)
    _protocol(got, ref64, ref32, mult=30.0)


def test_cfg4_full_size_vs_oracle(sb):
    """256 x 3 s crops, Fbank-80 + sentence MEAN normalisation."""
    rng = np.random.default_rng(704)
    lens = rng.uniform(0.5, 1.0, 256).astype(np.float32)
    lens[:3] = [1.0, 0.5, 32 * 5 / 301]
    wav = gc.synth_wav(704, 256, 48000, rel_lengths=lens)
    fb = sb.Fbank(n_mels=80)
    norm = sb.InputNormalization(norm_type="sentence", std_norm=False)
    feats = fb(dev(wav))
    got = feats.cpu().numpy()
    ref64 = _oracle_fbank(fb, wav, np.float64)
    ref32 = _oracle_fbank(fb, wav, np.float32)
    _protocol_full(got, ref64, ref32)
    out = norm(feats, torch.from_numpy(lens).cuda()).cpu().numpy()
    _check_sentence_norm(out, ref64, lens, std_norm=False)


def test_cfg5_shard_vs_oracle(sb):
    """One GPU's shard of cfg5: 128 x 30 s, Fbank-80 + global normalisation in training mode,
    two steps (first-call copy, then the running average).  Features of a slice of the shard
    against the float64 oracle; the normalisation of the WHOLE shard against the oracle
    normalising the product's own features (isolates the statistics / update / apply kernels)."""
    g = torch.Generator().manual_seed(705)
    fb = sb.Fbank(n_mels=80)
    gn = sb.InputNormalization(norm_type="global").train()
    ref = onp.InputNormalizationOracle(norm_type="global")
    for step in range(2):
        wav = torch.randn(128, 480000, generator=g) * 0.1
        lens = torch.linspace(0.6, 1.0, 128)[torch.randperm(128, generator=g)]
        for i in range(128):
            wav[i, int(round(float(lens[i]) * 480000)):] = 0
        feats = fb(wav.cuda())
        keep = feats.cpu().numpy()
        assert keep.shape == (128, 3001, 80)
        if step == 0:
            sl = [0, 17, 127]
            w = wav[sl].numpy()
            _protocol_full(keep[sl], _oracle_fbank(fb, w, np.float64), _oracle_fbank(fb, w, np.float32))
        out = gn(feats, lens.cuda(), epoch=0)
        assert torch.equal(feats.cpu(), torch.from_numpy(keep))          # x is left intact
        want = ref(keep.astype(np.float64), lens.numpy(), epoch=0)
        assert np.abs(out.cpu().numpy() - want).max() < 2e-4, step
        np.testing.assert_allclose(gn.glob_mean.cpu().numpy(), ref.glob_mean, rtol=1e-5, atol=2e-5)
        np.testing.assert_allclose(gn.glob_std.cpu().numpy(), ref.glob_std, rtol=2e-5, atol=2e-5)
    assert gn.count == ref.count == 2


@pytest.mark.parametrize("n_fft", [400, 512])
def test_sentence_norm_in_place_reads_before_it_writes(sb, n_fft):
    """In place, the CTAs of an utterance all READ the rows the statistics need (the tile the
    length cuts, tiles with values below the top_db floor) and one of them owns those rows in
    the normalisation pass: with more CTAs than the GPU holds at once (400 here) a CTA of a later
    wave must still see them un-normalised.  Ragged lengths, silence and near-silence make both
    kinds of rows; in place (after the call and beside its clamp launch) == out of place, bitwise,
    and == the reference's formula (processing/features.py:1020-1035) on the clamped features."""
    from speechbrain_b200 import ops

    torch.manual_seed(5)
    B, L = 200, 64000
    wav = torch.randn(B, L, device="cuda") * 0.1
    wav[::3, 20000:30000] = 0.0                       # digital silence: constant tiles
    wav[1::3, 40000:52000] *= 1e-6                    # below the floor: clamped columns
    lens = (0.3 + 0.7 * torch.rand(B, device="cuda")).clamp(max=1.0)
    lens[:4] = 1.0
    fb = sb.Fbank(n_mels=80, n_fft=n_fft)
    tab = fb._frontend_tables()
    for round_ in range(3):
        feats, _, ts = ops.fbank_forward(tab, wav, tile_stats=True, top_db=80.0)
        want = ops.norm_sentence_tiles(feats, lens, ts)
        keep = feats.clone()
        got = ops.norm_sentence_tiles(feats, lens, ts, out=feats)
        assert got.data_ptr() == feats.data_ptr()
        assert torch.equal(got, want), round_
        feats2, _, ts2 = ops.fbank_forward(tab, wav, tile_stats=True, top_db=80.0)
        got2 = ops.norm_sentence_tiles(feats2, lens, ts2, out=feats2, beside_clamp=True)
        assert torch.equal(got2, want), round_
    T = keep.shape[1]
    n = torch.round(lens * T).long().clamp(max=T)
    for b in (0, 1, 2, 3, 17, 101, 199):
        v = keep[b, : int(n[b])]
        ref = (keep[b] - v.mean(0)) / v.std(0).clamp(min=1e-10)
        assert torch.allclose(want[b], ref, rtol=1e-4, atol=2e-4), b


def test_normalisation_divides_exactly(sb):
    """The normalisation kernels give bitwise what the reference's `(x - mean) / std`
    (processing/features.py:1035) gives on the same GPU: a correctly rounded division - on 16 M
    random values with statistics over 12 binades, on all-ones mantissas, and where there is no
    finite quotient (std = 0, NaN statistics).  (A reciprocal + remainder + correction routine
    passes this test too and was measured: slower, the passes are not bound by the division.)"""
    from speechbrain_b200 import ops

    torch.manual_seed(11)
    B, T, F = 64, 3125, 80
    x = (torch.randn(B, T, F, device="cuda") * 30).contiguous()
    mean = torch.randn(B, F, device="cuda") * 10
    std = torch.exp(torch.randn(B, F, device="cuda") * 3).clamp(min=1e-10)
    std[0] = torch.nextafter(torch.full((F,), 2.0, device="cuda"), torch.zeros(F, device="cuda"))  # 1.11...1 x 2^0
    std[1, :8] = 0.0
    std[2, :8] = float("nan")
    got = ops.norm_apply(x, mean, std, per_utt=True)
    want = (x - mean[:, None, :]) / std[:, None, :]
    assert torch.equal(torch.nan_to_num(got, nan=7.0), torch.nan_to_num(want, nan=7.0))
    # the one-launch sentence kernel against the two-launch path's statistics
    wav = torch.randn(8, 48000, device="cuda") * 0.1
    lens = torch.linspace(0.5, 1.0, 8, device="cuda")
    fb = sb.Fbank(n_mels=80)
    feats, _, ts = ops.fbank_forward(fb._frontend_tables(), wav, tile_stats=True, top_db=80.0)
    one = ops.norm_sentence_tiles(feats, lens, ts)
    m, s = ops.norm_stats_tiles(feats, lens, ts)
    assert torch.equal(one, (feats - m[:, None, :]) / s[:, None, :])


def test_independent_batches_on_two_streams_match_the_single_stream_run(sb):
    """bench.py alternates consecutive batches between two streams (the normalisation of one
    batch then runs under the FFT kernel of the next).  Every step runs start to end on one
    stream, which is what the tile-record hand-over and SBFE_BESIDE_CLAMP need; eager and inside
    one CUDA graph the outputs are bitwise those of the single-stream run."""
    torch.manual_seed(9)
    B, L = 16, 96000
    lens = torch.linspace(0.55, 1.0, B, device="cuda")
    pool = []
    for i in range(6):
        w = torch.randn(B, L, device="cuda") * 0.1
        w[i % B, 30000:41000] = 0.0
        for b in range(B):
            w[b, int(round(float(lens[b]) * L)):] = 0
        pool.append(w)
    fb, norm = sb.Fbank(n_mels=80), sb.InputNormalization(norm_type="sentence")
    want = [norm(fb(w), lens).clone() for w in pool]
    torch.cuda.synchronize()
    lanes = [torch.cuda.Stream(), torch.cuda.Stream()]
    for s in lanes:
        s.wait_stream(torch.cuda.current_stream())
    got = []
    for i, w in enumerate(pool):                      # eager
        with torch.cuda.stream(lanes[i % 2]):
            got.append(norm(fb(w), lens))
    torch.cuda.synchronize()
    for a, b in zip(got, want):
        assert torch.equal(a, b)
    g, keep = torch.cuda.CUDAGraph(), []
    with torch.cuda.graph(g):                         # one graph, two parallel chains
        main = torch.cuda.current_stream()
        side = lanes[0]
        fork = torch.cuda.Event()
        fork.record(main)
        side.wait_event(fork)
        for i, w in enumerate(pool):
            with torch.cuda.stream(main if i % 2 == 0 else side):
                keep.append(norm(fb(w), lens))
        join = torch.cuda.Event()
        join.record(side)
        main.wait_event(join)
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    for a, b in zip(keep, want):
        assert torch.equal(a, b)
