"""world_size-2 gloo tests (CPU) of the one collective on the path: the [2F+1] statistics
payload all-reduce of InputNormalization(sync_stats=True).

The PRODUCT's InputNormalization.forward runs in both ranks (its host logic: mode dispatch,
count / weight / epoch gating, payload sync) with CPU stand-ins for the CUDA kernels
(tests/dist_workers.py: install_cpu_kernels); every rank's output must equal the corresponding
slice of the ORACLE run single-process on the concatenated global batch - which is how the
reference itself would behave on that batch (processing/features.py:1083-1114) - with uneven
shards and the running average over three steps."""
import os
import sys

import numpy as np
import pytest
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import dist_workers as dw  # noqa: E402


def _oracle_on_concatenation(norm_type, world):
    from oracle import frontend_np as onp

    ref = onp.InputNormalizationOracle(norm_type=norm_type)
    outs = []
    for step in range(dw.STEPS):
        xs, ls = zip(*[dw.feat_shard(r, step) for r in range(world)])
        outs.append(ref(np.concatenate(xs).copy(), np.concatenate(ls), epoch=0))
    return outs, ref


@pytest.mark.parametrize("norm_type", ["global", "batch"])
def test_synced_statistics_equal_the_concatenated_batch(norm_type):
    world = 2
    port = 29500 + (os.getpid() % 2000) + (7 if norm_type == "batch" else 0)
    out = mp.Manager().dict()
    mp.spawn(dw.cpu_worker, args=(world, port, out, norm_type), nprocs=world, join=True)
    want, ref = _oracle_on_concatenation(norm_type, world)
    for step in range(dw.STEPS):
        lo = 0
        for r in range(world):
            got = out[r][0][step]
            np.testing.assert_allclose(got, want[step][lo:lo + got.shape[0]], rtol=2e-5, atol=2e-5)
            lo += got.shape[0]
    if norm_type == "global":
        for r in range(world):
            np.testing.assert_allclose(out[r][1], ref.glob_mean, rtol=1e-5, atol=1e-5)
            assert out[r][2] == ref.count == dw.STEPS
        # both ranks hold identical running statistics afterwards
        np.testing.assert_array_equal(out[0][1], out[1][1])


def test_sync_payload_is_a_no_op_without_a_process_group():
    import torch

    from speechbrain_b200.processing.normalization import sync_payload

    cur = torch.arange(5.0)
    assert sync_payload(cur) is cur and torch.equal(cur, torch.arange(5.0))
