"""CPU-side checks: the C-ABI library builds/loads and exports every symbol that
include/sbfe.h declares, host tables match the reference's constants, and the
product refuses CPU tensors instead of falling back."""
import os
import re
import subprocess

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    from speechbrain_b200 import _lib

    handle = _lib.lib()
    header = open(os.path.join(ROOT, "include", "sbfe.h")).read()
    declared = set(re.findall(r"SBFE_API [\w\*\s]+?(sbfe_\w+)\(", header))
    assert declared and declared == set(_lib.SIGNATURES)
    for name in declared:
        assert getattr(handle, name) is not None
    assert handle.sbfe_abi_version() == _lib.ABI_VERSION == 14
    assert handle.sbfe_num_frames(16000, 160) == 101
    assert handle.sbfe_error_string(-1).decode().startswith("sbfe")


def test_library_is_sm100a_and_has_no_oracle_dependency():
    from speechbrain_b200 import _build

    out = subprocess.run(["cuobjdump", "-lelf", _build.LIB], capture_output=True, text=True).stdout
    assert "sm_100a" in out
    for dirpath, _, files in os.walk(os.path.join(ROOT, "speechbrain_b200")):
        for f in files:
            if f.endswith(".py"):
                src = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in src, "product must not touch the oracle: " + f


def test_host_codelets_match_naive_dft(tmp_path):
    for name in ("codelet_check", "frame400_check", "frame400_v3_check"):
        exe = tmp_path / name
        subprocess.run(["g++", "-O2", "-o", str(exe), os.path.join(ROOT, "tests", "host", name + ".cpp")],
                       check=True)
        res = subprocess.run([str(exe)], capture_output=True, text=True)
        assert res.returncode == 0, res.stdout


def test_host_tables_match_reference_constants(golden):
    import speechbrain_b200 as sb

    fb = sb.Fbank(n_mels=80)
    name = "cfg2_fbank80_sentnorm_4x2s"
    assert np.array_equal(fb.compute_fbanks.f_central.numpy(), golden[f"{name}/fc"])
    assert np.array_equal(fb.compute_fbanks.band.numpy(), golden[f"{name}/band"])
    assert np.array_equal(fb.compute_STFT.window.numpy(), golden[f"{name}/window"])
    tab = fb._frontend_tables()
    dense = fb.compute_fbanks.fbank_matrix()
    start, count, offset, w = tab._h_mel
    assert int(count.sum()) == 387 and tab.n_mels == 80  # 387 non-zeros for 80 mels
    # padded layout of sbfe_fbank_desc.mel_pad4 == 2: offsets multiples of 4, zero fill, filters
    # (2k, 2k+1) store the same number of 4-weight groups, start + 4 G stays inside 204 rows
    assert all(int(o) % 4 == 0 for o in offset) and tab.nnz == w.numel() and tab.nnz % 4 == 0
    stored = torch.diff(torch.cat([offset, torch.tensor([tab.nnz], dtype=torch.int32)]))
    assert torch.equal(stored[0::2], stored[1::2])
    assert bool((stored >= 4 * ((count + 3) // 4)).all()) and int((start + stored).max()) <= 204
    assert int(torch.count_nonzero(w)) == 387
    rebuilt = torch.zeros_like(dense)
    for m in range(80):
        s, c, o = int(start[m]), int(count[m]), int(offset[m])
        rebuilt[s:s + c, m] = w[o:o + c]
    assert torch.equal(rebuilt, dense)
    assert tab.floor_db == -100.0


def test_drop_in_surface():
    import speechbrain_b200 as sb

    assert list(sb.Fbank().state_dict()) == ["compute_deltas.kernel"]
    assert list(sb.MFCC().state_dict()) == ["compute_deltas.kernel"]
    assert sb.MFCC().state_dict()["compute_deltas.kernel"].shape == (20, 1, 5)
    n = sb.InputNormalization()
    assert set(n._statistics_dict()) == {"count", "glob_mean", "glob_std", "spk_dict_mean",
                                         "spk_dict_std", "spk_dict_count"}
    import inspect
    assert list(inspect.signature(n.forward).parameters)[:2] == ["x", "lengths"]
    with pytest.raises(ValueError):
        sb.DCT(input_size=10, n_out=20)
    assert not sb.Fbank(requires_grad=True).compute_fbanks.freeze   # learnable: tests/test_learnable.py


def test_cpu_tensors_are_refused():
    import speechbrain_b200 as sb

    with pytest.raises(RuntimeError, match="CUDA only"):
        sb.Fbank()(torch.randn(2, 1600))
    with pytest.raises(RuntimeError, match="CUDA only"):
        sb.InputNormalization()(torch.randn(2, 10, 4), torch.ones(2))


def test_checkpoint_roundtrip(tmp_path):
    import speechbrain_b200 as sb

    a = sb.InputNormalization()
    a.count, a.glob_mean, a.glob_std = 7, torch.arange(4.0), torch.ones(4) * 2
    a.spk_dict_mean[3], a.spk_dict_std[3], a.spk_dict_count[3] = torch.zeros(4), torch.ones(4), 2
    p = str(tmp_path / "normalizer.ckpt")
    a._save(p)
    raw = torch.load(p)
    assert set(raw) == {"count", "glob_mean", "glob_std", "spk_dict_mean", "spk_dict_std", "spk_dict_count"}
    b = sb.InputNormalization()
    b._load(p, end_of_epoch=False, device="cpu")
    assert b.count == 7 and torch.equal(b.glob_mean, a.glob_mean) and b.spk_dict_count == {3: 2}
    # legacy checkpoints hold python ints (reference features.py:1169-1171)
    torch.save({**raw, "glob_mean": 0, "glob_std": 0}, p)
    b._load(p)
    assert b.glob_mean == 0


def test_jit_trace_bridge_records_one_operator():
    """speechbrain_b200/_jit.py: under torch.jit.trace a module forward becomes one
    sbfe::module_call node that re-enters the Python forward on replay (no kernels needed
    to check the mechanism: the body below is invisible to the tracer, like a ctypes call)."""
    import torch

    from speechbrain_b200 import _jit

    class Twice(torch.nn.Module):
        @_jit.traceable()
        def forward(self, x):
            return torch.from_numpy(x.numpy() * 2)

    class MinusInPlace(torch.nn.Module):
        @_jit.traceable(0)
        def forward(self, x, lengths, epoch=0):
            x.copy_(torch.from_numpy(x.numpy() - 1 - epoch))
            return x

    t = torch.jit.trace(Twice(), (torch.ones(3),))
    assert "sbfe::module_call" in str(t.graph)
    assert torch.equal(t(torch.full((3,), 5.0)), torch.full((3,), 10.0))
    m = MinusInPlace()
    t = torch.jit.trace(m, (torch.ones(3), torch.ones(1)))
    y = torch.full((3,), 5.0)
    r = t(y, torch.ones(1))
    assert torch.equal(r, torch.full((3,), 4.0)) and torch.equal(y, r)
    # eager calls are untouched by the decorator
    assert torch.equal(m(torch.full((3,), 5.0), torch.ones(1), epoch=2), torch.full((3,), 2.0))


def test_mel_layout_moves_a_filter_that_would_leave_the_spectrum():
    """mel_pad4 == 2 (include/sbfe.h): a narrow filter at the top of the spectrum paired with a
    wide one stores the pair's group count; it is moved down with leading zeros so that
    start + 4 G <= 204, and the table still describes the same matrix."""
    import torch

    from speechbrain_b200 import ops

    mat = torch.zeros(201, 3)
    mat[180:201, 0] = torch.arange(1, 22).float()
    mat[199:201, 1] = torch.tensor([0.5, 0.25])
    mat[5:7, 2] = 1.0
    tab = ops.FrontendTables(400, 160, torch.hamming_window(400), mat)
    start, count, offset, w = tab._h_mel
    stored = torch.diff(torch.cat([offset, torch.tensor([tab.nnz], dtype=torch.int32)]))
    assert stored.tolist() == [24, 24, 4]
    assert int(start[1]) == 180 and int(count[1]) == 21 and int((start + stored).max()) <= 204
    rebuilt = torch.zeros_like(mat)
    for m in range(3):
        s, c, o = int(start[m]), int(count[m]), int(offset[m])
        rebuilt[s:s + c, m] = w[o:o + c]
        assert torch.count_nonzero(w[o + c:o + int(stored[m])]) == 0
    assert torch.equal(rebuilt, mat)


def test_pair_is_a_harmless_shim():
    """pair() (lobes/features.py) survives for round-1 recipes: a type-checked no-op, now that the
    fused kernel finishes the clamp itself."""
    import speechbrain_b200 as sb

    fb, norm = sb.Fbank(n_mels=80), sb.InputNormalization(norm_type="sentence")
    assert sb.pair(fb, norm) is fb and sb.pair(fb, None) is fb
    with pytest.raises(TypeError):
        sb.pair(torch.nn.Identity(), norm)
    assert sb.pair(sb.MFCC(), norm) is not None
