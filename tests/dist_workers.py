"""Workers for the multi-rank tests of the path's one collective: the [2F+1]
statistics all-reduce of ``InputNormalization(norm_type="global"|"batch", sync_stats=True)``.

Spawned with ``torch.multiprocessing.spawn`` by tests/test_dist_cpu.py (gloo, CPU: the product's
host logic with CPU stand-ins for the kernels) and tests/test_dist_gpu.py (the real kernels:
NCCL with one rank per GPU when the box has enough GPUs, else both ranks on cuda:0 over gloo,
which stages CUDA tensors through the host - no kernel ever waits for another rank).

The reference under DDP does NOT synchronise these statistics (core.py:1243-1244: every rank
keeps its own); the definition of parity for the sync is "equals the reference run in one
process on the concatenated global batch" (processing/features.py:1083-1114), which is what the
callers check against the oracle.
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)

STEPS = 3
F_CPU = 12


def feat_shard(rank, step):
    """Uneven feature shards for the CPU test: rank r holds 3 + 2r sentences."""
    rng = np.random.default_rng(900 + 10 * step + rank)
    b = 3 + 2 * rank
    x = (rng.standard_normal((b, 41, F_CPU)) * 3 - 20).astype(np.float32)
    lens = rng.uniform(0.3, 1.0, size=b).astype(np.float32)
    lens[0] = 1.0
    return x, lens


def wav_shard(rank, step, B=5, L=2 * 16000 + 40):
    """Uneven waveform shards for the GPU test: rank r holds B + r utterances."""
    rng = np.random.default_rng(7000 + 10 * step + rank)
    n = B + rank
    wav = (rng.standard_normal((n, L)) * 0.1).astype(np.float32)
    lens = (rng.uniform(0.5, 1.0, size=n)).astype(np.float32)
    lens[0] = 1.0
    for i in range(n):
        wav[i, int(round(float(lens[i]) * L)):] = 0
    return wav, lens


def install_cpu_kernels():
    """CPU stand-ins for the CUDA entry points InputNormalization.forward calls, so that the
    PRODUCT's host logic (count, weight, epoch gating, payload sync, mode dispatch) runs in a
    CPU process.  Test infrastructure only: the stand-ins restate include/sbfe.h in numpy."""
    from oracle import frontend_np as onp
    from speechbrain_b200 import ops

    def norm_stats(x, lengths, mean_norm=True, std_norm=True, eps=1e-10, utt_max=None, top_db=80.0):
        m, s = onp.sentence_stats(x.numpy(), lengths.numpy(), mean_norm, std_norm, eps)
        f = x.shape[2]
        return (torch.from_numpy(np.broadcast_to(m, (x.shape[0], f)).copy()),
                torch.from_numpy(np.broadcast_to(s, (x.shape[0], f)).copy()))

    def norm_reduce(mean, std):
        return torch.cat([mean.sum(0), std.sum(0), torch.tensor([float(mean.shape[0])])])

    def norm_update(cur, gm, gs, first, weight):
        f = gm.numel()
        n = cur[2 * f]
        if first:
            gm.copy_(cur[:f] / n)
            gs.copy_(cur[f:2 * f] / n)
        else:
            gm.copy_((1 - weight) * gm + weight * (cur[:f] / n))
            gs.copy_((1 - weight) * gs + weight * (cur[f:2 * f] / n))

    def norm_apply(x, mean, std, per_utt, out=None, utt_max=None, top_db=80.0, out_dtype=None):
        y = (x - mean[:, None, :]) / std[:, None, :] if per_utt else (x - mean) / std
        if out is None:
            return y
        out.copy_(y)
        return out

    ops._require_cuda = lambda t, name: None
    ops.norm_stats, ops.norm_reduce = norm_stats, norm_reduce
    ops.norm_update, ops.norm_apply = norm_update, norm_apply


def cpu_worker(rank, world, port, out, norm_type):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.distributed.init_process_group("gloo", rank=rank, world_size=world)
    try:
        install_cpu_kernels()
        import speechbrain_b200 as sb

        norm = sb.InputNormalization(norm_type=norm_type, sync_stats=True).train()
        res = []
        for step in range(STEPS):
            x, lens = feat_shard(rank, step)
            y = norm(torch.from_numpy(x.copy()), torch.from_numpy(lens), epoch=0)
            res.append(y.numpy().copy())
        gm = norm.glob_mean.numpy().copy() if norm_type == "global" else None
        out[rank] = (res, gm, norm.count)
    finally:
        torch.distributed.destroy_process_group()


def gpu_worker(rank, world, port, out, backend):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dev = torch.device("cuda", rank if backend == "nccl" else 0)
    torch.cuda.set_device(dev)
    kw = {"device_id": dev} if backend == "nccl" else {}
    torch.distributed.init_process_group(backend, rank=rank, world_size=world, **kw)
    try:
        import speechbrain_b200 as sb

        fb = sb.Fbank(n_mels=80)
        norm = sb.InputNormalization(norm_type="global", sync_stats=True).train()
        res = []
        for step in range(STEPS):
            wav, lens = wav_shard(rank, step)
            y = norm(fb(torch.from_numpy(wav).to(dev)), torch.from_numpy(lens).to(dev), epoch=0)
            res.append(y.cpu().numpy())
        stats = torch.cat([norm.glob_mean, norm.glob_std])
        ref0 = stats.clone()
        torch.distributed.broadcast(ref0, 0)
        same = bool(torch.equal(stats, ref0))
        out[rank] = (res, norm.glob_mean.cpu().numpy(), norm.glob_std.cpu().numpy(), norm.count, same)
    finally:
        torch.distributed.destroy_process_group()
