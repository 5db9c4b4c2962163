"""sbfe_norm_stats_static: the no-allocation variant of sbfe_norm_stats (include/sbfe.h).

No counterpart in the reference: the partial sums
live in a device array built into libsbfe.so instead of a caller-provided workspace.  Called
straight through the C ABI, checked against the existing entry point on two calls in a row
(the array keeps the last call's contents).  Collected last on purpose (file name).
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

STATIC_WS_BYTES = 2 * 1024 * 1024   # SBFE_NORM_STATIC_WS_BYTES
E_STATIC_LIMIT = -4


def _both(L_, x, lens, utt_max=None, top_db=80.0):
    """(mean, std) from sbfe_norm_stats (caller workspace) and from sbfe_norm_stats_static."""
    B, T, F = x.shape
    ptr = lambda t: None if t is None else t.data_ptr()  # noqa: E731
    stream = torch.cuda.current_stream().cuda_stream
    ws_bytes = L_.sbfe_norm_workspace_bytes(B, T, F)
    ws = torch.empty((max(ws_bytes, 16),), dtype=torch.uint8, device=x.device)
    m0, s0, m1, s1 = (torch.full((B, F), float("nan"), device=x.device) for _ in range(4))
    rc = L_.sbfe_norm_stats(ptr(x), ptr(lens), B, T, F, 1, 1, 1e-10, ptr(utt_max), top_db,
                            ptr(m0), ptr(s0), ptr(ws), ws_bytes, stream)
    assert rc == 0, L_.sbfe_error_string(rc)
    rc = L_.sbfe_norm_stats_static(ptr(x), ptr(lens), B, T, F, 1, 1, 1e-10, ptr(utt_max), top_db,
                                   ptr(m1), ptr(s1), stream)
    assert rc == 0, L_.sbfe_error_string(rc)
    torch.cuda.synchronize()
    return (m0, s0), (m1, s1)


def test_static_scratch_entry_point_matches_the_workspace_one_twice_in_a_row():
    from speechbrain_b200 import _lib

    L_ = _lib.lib()
    g = torch.Generator(device="cpu").manual_seed(11)
    # two calls in a row with different shapes: the second one finds the first one's partial sums
    # in the shared array (more chunks, then fewer) and must not be affected by them
    for B, T, F, shift in ((6, 1601, 80, 3.0), (9, 300, 40, -7.0)):
        x = (torch.randn(B, T, F, generator=g) * 4.0 + shift).cuda()
        lens = torch.linspace(0.3, 1.0, B).cuda()
        (m0, s0), (m1, s1) = _both(L_, x, lens)
        # same kernels, same order of operations: the tolerance of the norm_stats tests is 2e-5
        assert torch.isfinite(m1).all() and torch.isfinite(s1).all()
        assert float((m1 - m0).abs().max()) <= 2e-5 and float((s1 - s0).abs().max()) <= 2e-5
        assert torch.equal(m1, m0) and torch.equal(s1, s0)
        n = torch.round(lens * T).int().tolist()
        want = torch.stack([x[b, :n[b]].double().mean(0) for b in range(B)]).float()
        assert float((m1 - want).abs().max()) <= 2e-5 * max(1.0, float(want.abs().max()))


def test_static_scratch_entry_point_rejects_what_does_not_fit():
    from speechbrain_b200 import _lib

    L_ = _lib.lib()
    B, T, F = 64, 8192, 80                       # 16 * 64 * 32 * 80 = 2.5 MiB > 2 MiB
    assert L_.sbfe_norm_workspace_bytes(B, T, F) > STATIC_WS_BYTES
    x = torch.zeros(1, device="cuda")            # never dereferenced: the call must refuse first
    rc = L_.sbfe_norm_stats_static(x.data_ptr(), x.data_ptr(), B, T, F, 1, 1, 1e-10, None, 80.0,
                                   x.data_ptr(), x.data_ptr(), torch.cuda.current_stream().cuda_stream)
    assert rc == E_STATIC_LIMIT
    assert b"SBFE_NORM_STATIC_WS_BYTES" in L_.sbfe_error_string(rc)
    torch.cuda.synchronize()
    # the largest shape of the path's configs that fits: one cfg5 shard column block
    assert L_.sbfe_norm_workspace_bytes(128, 3001, 80) <= STATIC_WS_BYTES
