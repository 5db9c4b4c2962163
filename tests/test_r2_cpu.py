"""Round-2 CPU checks: the oracles against the round-2 reference fixtures (filterbank jitter,
other hops, n_fft 512, gaussian filters, normalized STFT), the timed torch-CPU port
(oracle/frontend_torch.py - the denominator of every reported ratio) against the goldens,
reference-written ``normalizer.ckpt`` files, and ``install()`` + checkpoint hooks against the
reference package when it is present (this container; not the GPU box)."""
import os
import sys
import types

import numpy as np
import pytest
import torch

import golden_cases as gc
from oracle import frontend_np as onp

REF_ROOT = os.environ.get("SB_REFERENCE", "/root/reference")
LOBES_R2 = [n for n, s in gc.CASES_R2.items() if s["kind"] in ("fbank", "mfcc")]


@pytest.fixture(scope="module")
def golden2():
    return np.load(gc.GOLDEN_R2_NPZ)


def oracle_lobe(spec, x, dtype, g, name):
    kw = dict(spec["kwargs"])
    kw.pop("param_rand_factor", None)
    fc, band = g[f"{name}/fc"], g[f"{name}/band"]
    if "train_seed" in spec:   # processing/features.py:525-532: both scaled by the same draw
        r = g[f"{name}/rand"]
        jit = 1.0 + r * 2 * spec["kwargs"]["param_rand_factor"] - spec["kwargs"]["param_rand_factor"]
        fc, band = (fc * np.float32(jit[0])).astype(np.float32), (band * np.float32(jit[1])).astype(np.float32)
    kw["mel_params"], kw["window"] = (fc, band), g[f"{name}/window"]
    if "deltas" in kw:
        kw["deltas_"] = kw.pop("deltas")
    fn = onp.mfcc if spec["kind"] == "mfcc" else onp.fbank
    return fn(x, dtype=dtype, **kw)


@pytest.mark.parametrize("name", LOBES_R2)
def test_oracle_matches_reference_r2(golden2, name):
    spec = gc.CASES_R2[name]
    x = spec["make_input"]()
    np.testing.assert_allclose(gc.input_checksum(x), golden2[f"{name}/in_ck"], rtol=1e-12)
    ref32 = golden2[f"{name}/ref32"]
    got32 = oracle_lobe(spec, x, np.float32, golden2, name)
    assert got32.shape == ref32.shape
    mult = 30.0 if spec["kind"] == "mfcc" else 1.0
    # gaussian filters sum all 201 bins in float32: the summation order differs (GEMM blocking)
    loose = 5.0 if spec["kwargs"].get("filter_shape") == "gaussian" else 1.0
    assert np.abs(got32 - ref32).max() < 2e-3 * mult * loose
    assert np.abs(got32 - ref32).mean() < 2e-5 * mult * loose
    if f"{name}/ref64" in golden2.files:
        got64 = oracle_lobe(spec, x, np.float64, golden2, name)
        assert np.abs(got64 - golden2[f"{name}/ref64"]).max() < 2e-5 * mult


def test_jitter_draw_is_the_default_generator(golden2):
    """The fixture's jitter factors are what torch.rand(2) gives after manual_seed: the product
    makes the same call in the same place (processing/features.py: Filterbank.tables)."""
    for name in ("fbank_jitter_train", "mfcc_jitter_train"):
        torch.manual_seed(gc.CASES_R2[name]["train_seed"])
        assert np.array_equal(torch.rand(2).numpy(), golden2[f"{name}/rand"])
    import speechbrain_b200 as sb

    fb = sb.Filterbank(n_mels=40, param_rand_factor=0.15).train()
    torch.manual_seed(gc.JITTER_SEED)
    tab = fb.tables()
    r = golden2["fbank_jitter_train/rand"]
    fc = fb.f_central * float(1.0 + r[0] * 0.3 - 0.15)
    band = fb.band * float(1.0 + r[1] * 0.3 - 0.15)
    want = onp.filterbank_matrix(fc.numpy(), band.numpy(), np.linspace(0, 8000, 201, dtype=np.float32))
    start, count, offset, w = tab._h_mel
    dense = np.zeros((201, 40), np.float32)
    for m in range(40):
        s, c, o = int(start[m]), int(count[m]), int(offset[m])
        dense[s:s + c, m] = w[o:o + c].numpy()
    assert np.abs(dense - want).max() < 1e-5
    assert fb.eval().tables() is fb.tables()       # eval mode: cached, no draw


def test_stft_oracle_matches_reference_r2(golden2):
    for name in ("stft_normalized", "stft_normalized_nfft512"):
        spec = gc.CASES_R2[name]
        x = spec["make_input"]()
        kw = dict(spec["kwargs"])
        assert kw.pop("normalized_stft")
        n_fft = kw.get("n_fft", 400)
        got = onp.stft(x, **kw) * np.float32(n_fft ** -0.5)   # torch.stft(normalized=True)
        ref = golden2[f"{name}/ref32"]
        assert got.shape == ref.shape and np.abs(got - ref).max() < 2e-5 * max(1.0, np.abs(ref).max())


def test_torch_cpu_port_is_pinned(golden):
    """oracle/frontend_torch.py (timed by bench.py as the CPU baseline and the reference arm):
    bit-identical to the unmodified reference on the cfg2 / cfg4 goldens."""
    from oracle import frontend_torch as ot

    for name in ("cfg2_fbank80_sentnorm_4x2s", "cfg4_fbank80_sentmean_4x1s"):
        spec = gc.CASES[name]
        x = torch.from_numpy(spec["make_input"]())
        torch.set_num_threads(1)
        feats = ot.fbank(x, n_mels=80)
        assert np.array_equal(feats.numpy(), golden[f"{name}/ref32"])
        lens = torch.tensor(spec["lengths"], dtype=torch.float32)
        out = ot.sentence_norm(feats.clone(), lens, std_norm=spec["norm"].get("std_norm", True))
        assert np.array_equal(out.numpy(), golden[f"{name}/norm32"])
    x = torch.from_numpy(gc.CASES["cfg2_fbank80_sentnorm_4x2s"]["make_input"]())
    lens = torch.tensor(gc.LENS4, dtype=torch.float32)
    assert np.array_equal(ot.fbank_sentence_norm(x, lens, n_mels=80).numpy(),
                          golden["cfg2_fbank80_sentnorm_4x2s/norm32"])
    assert np.array_equal(ot.fbank(torch.from_numpy(gc.CASES["cfg1_fbank40_8x1s"]["make_input"]())).numpy(),
                          golden["cfg1_fbank40_8x1s/ref32"])


@pytest.mark.parametrize("norm_type,path", [("global", gc.REF_CKPT_GLOBAL), ("speaker", gc.REF_CKPT_SPEAKER)])
def test_product_loads_reference_written_checkpoint(norm_type, path, golden2, tmp_path):
    """normalizer.ckpt written by the REFERENCE's InputNormalization._save (features.py:1205-1215)
    -> product _load: same state; product _save -> the same six keys and values."""
    import speechbrain_b200 as sb

    raw = torch.load(path)
    norm = sb.InputNormalization(norm_type=norm_type)
    norm._load(path, end_of_epoch=False, device="cpu")
    assert norm.count == raw["count"] == int(golden2[f"ckpt_{norm_type}/count"])
    assert torch.equal(torch.as_tensor(norm.glob_mean), torch.as_tensor(raw["glob_mean"]))
    assert set(norm.spk_dict_mean) == set(raw["spk_dict_mean"])
    for k in raw["spk_dict_mean"]:
        assert torch.equal(norm.spk_dict_mean[k], raw["spk_dict_mean"][k])
        assert torch.equal(norm.spk_dict_std[k], raw["spk_dict_std"][k])
    assert norm.spk_dict_count == raw["spk_dict_count"]
    p2 = str(tmp_path / "normalizer.ckpt")
    norm._save(p2)
    again = torch.load(p2)
    assert set(again) == set(raw) and again["count"] == raw["count"]
    assert torch.equal(torch.as_tensor(again["glob_std"]), torch.as_tensor(raw["glob_std"]))
    assert again["spk_dict_count"] == raw["spk_dict_count"]


def _reference_shells():
    ref = os.path.join(REF_ROOT, "speechbrain")
    if not os.path.isdir(ref):
        pytest.skip("the reference checkout is not on this machine")
    saved = {k: v for k, v in sys.modules.items() if k == "speechbrain" or k.startswith("speechbrain.")}

    def shell(name, path):
        m = types.ModuleType(name)
        m.__path__ = [path]
        sys.modules[name] = m

    shell("speechbrain", ref)
    for sub in ("utils", "processing", "nnet", "lobes", "dataio"):
        shell("speechbrain." + sub, os.path.join(ref, sub))
    return saved


def test_install_rebinds_the_reference_names_and_registers_checkpoint_hooks(tmp_path):
    """install() against the real reference modules: YAML's dotted paths then resolve to the
    B200 classes, and the reference Checkpointer's hook lookup (utils/checkpoints.py:330-359)
    finds _save / _load for InputNormalization, so a reference Checkpointer round-trips the
    statistics instead of saving an empty state_dict."""
    saved = _reference_shells()
    try:
        import speechbrain.lobes.features as lobes
        import speechbrain.processing.features as proc
        from speechbrain.utils import checkpoints as ck

        ref_fbank, ref_norm = lobes.Fbank, proc.InputNormalization
        import speechbrain_b200 as sb

        # without install(): constructing the class registers the hooks once the application
        # has imported the reference's checkpoints module
        for table in (ck.DEFAULT_SAVE_HOOKS, ck.DEFAULT_LOAD_HOOKS, ck.DEFAULT_TRANSFER_HOOKS):
            table.pop(sb.InputNormalization, None)
        assert ck.get_default_hook(sb.InputNormalization(), ck.DEFAULT_SAVE_HOOKS) == sb.InputNormalization._save
        patched = sb.install()
        assert "speechbrain.lobes.features.Fbank" in patched
        assert lobes.Fbank is sb.Fbank and lobes.MFCC is sb.MFCC
        assert proc.InputNormalization is sb.InputNormalization and proc.STFT is sb.STFT
        assert lobes.Fbank is not ref_fbank and proc.InputNormalization is not ref_norm
        norm = sb.InputNormalization()
        assert ck.get_default_hook(norm, ck.DEFAULT_SAVE_HOOKS) == sb.InputNormalization._save
        assert ck.get_default_hook(norm, ck.DEFAULT_LOAD_HOOKS) == sb.InputNormalization._load
        assert ck.get_default_hook(norm, ck.DEFAULT_TRANSFER_HOOKS) == sb.InputNormalization._load
        # a reference Checkpointer saves and recovers the running statistics
        norm.count, norm.glob_mean, norm.glob_std = 5, torch.arange(4.0), torch.ones(4) * 3
        cp = ck.Checkpointer(str(tmp_path), recoverables={"normalizer": norm})
        cp.save_checkpoint()
        fresh = sb.InputNormalization()
        cp2 = ck.Checkpointer(str(tmp_path), recoverables={"normalizer": fresh})
        assert cp2.recover_if_possible() is not None
        assert fresh.count == 5 and torch.equal(fresh.glob_mean, torch.arange(4.0))
    finally:
        for k in [k for k in sys.modules if k == "speechbrain" or k.startswith("speechbrain.")]:
            del sys.modules[k]
        sys.modules.update(saved)
