"""world_size-2 gloo test (CPU) of the one collective on the path: the [2F+1] statistics
payload all-reduce of InputNormalization(sync_stats=True).  Each rank builds its payload with
the numpy oracle from its own (uneven) shard, the product's sync_payload() all-reduces it, and
the resulting global mean / std must equal the oracle run on the concatenated batch
("equals the reference run single-process on the concatenated global batch")."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

SHAPES = [(3, 41, 12), (5, 41, 12)]  # uneven shards


def _shard(rank):
    rng = np.random.default_rng(900 + rank)
    b, t, f = SHAPES[rank]
    x = (rng.standard_normal((b, t, f)) * 3 - 20).astype(np.float32)
    lens = rng.uniform(0.3, 1.0, size=b).astype(np.float32)
    lens[0] = 1.0
    return x, lens


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.distributed.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import frontend_np as onp
        from speechbrain_b200.processing.normalization import sync_payload

        x, lens = _shard(rank)
        mean, std = onp.sentence_stats(x, lens)  # [B, F] each
        f = x.shape[2]
        cur = torch.zeros(2 * f + 1)
        cur[:f] = torch.from_numpy(mean.sum(0))
        cur[f:2 * f] = torch.from_numpy(std.sum(0))
        cur[2 * f] = x.shape[0]
        cur = sync_payload(cur)
        n = float(cur[2 * f])
        out[rank] = (cur[:f] / n).numpy().copy(), (cur[f:2 * f] / n).numpy().copy(), n
    finally:
        torch.distributed.destroy_process_group()


def test_sync_payload_equals_concatenated_batch():
    from oracle import frontend_np as onp

    if not hasattr(onp, "sentence_stats"):
        pytest.skip("oracle has no sentence_stats")
    port = 29500 + (os.getpid() % 2000)
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    xs, ls = zip(*[_shard(r) for r in range(2)])
    mean, std = onp.sentence_stats(np.concatenate(xs), np.concatenate(ls))
    want_m, want_s = mean.mean(0), std.mean(0)
    for r in range(2):
        got_m, got_s, n = out[r]
        assert n == 8.0
        np.testing.assert_allclose(got_m, want_m, rtol=1e-5, atol=1e-5)
        np.testing.assert_allclose(got_s, want_s, rtol=1e-5, atol=1e-5)
    # both ranks hold identical statistics afterwards
    np.testing.assert_array_equal(out[0][0], out[1][0])
    np.testing.assert_array_equal(out[0][1], out[1][1])
