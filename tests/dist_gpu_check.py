"""Multi-GPU parity of the one collective: run under torchrun with N >= 2 ranks (one per GPU).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29611 tests/dist_gpu_check.py

Every rank computes Fbank(80) + InputNormalization("global", sync_stats=True) on ITS shard for
three training steps (NCCL all-reduce of the [2F+1] payload) and, separately, the same modules
single-process (sync_stats=False) on the concatenation of all ranks' shards.  The synced result
must equal the corresponding slice of the concatenated run, and the running statistics must be
identical on every rank."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import speechbrain_b200 as sb  # noqa: E402


def shard(rank, step, B=6, L=3 * 16000 + 40):
    g = torch.Generator().manual_seed(7000 + 10 * step + rank)
    wav = torch.randn(B + rank, L, generator=g) * 0.1  # uneven shards
    lens = torch.rand(B + rank, generator=g) * 0.5 + 0.5
    lens[0] = 1.0
    for i in range(wav.shape[0]):
        wav[i, int(round(float(lens[i]) * L)):] = 0
    return wav, lens


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    torch.distributed.init_process_group("nccl", device_id=dev)
    fb = sb.Fbank(n_mels=80)
    n_sync = sb.InputNormalization(norm_type="global", sync_stats=True).train()
    n_ref = sb.InputNormalization(norm_type="global", sync_stats=False).train()
    worst = 0.0
    for step in range(3):
        wav, lens = shard(rank, step)
        got = n_sync(fb(wav.to(dev)), lens.to(dev), epoch=0)
        parts = [shard(r, step) for r in range(world)]
        cat_w = torch.cat([p[0] for p in parts]).to(dev)
        cat_l = torch.cat([p[1] for p in parts]).to(dev)
        want = n_ref(fb(cat_w), cat_l, epoch=0)
        lo = sum(p[0].shape[0] for p in parts[:rank])
        err = float((got - want[lo:lo + wav.shape[0]]).abs().max())
        worst = max(worst, err)
        assert err < 2e-4, (rank, step, err)
        assert torch.allclose(n_sync.glob_mean, n_ref.glob_mean, rtol=1e-5, atol=1e-5)
        assert torch.allclose(n_sync.glob_std, n_ref.glob_std, rtol=1e-5, atol=1e-5)
    stats = torch.cat([n_sync.glob_mean, n_sync.glob_std])
    ref0 = stats.clone()
    torch.distributed.broadcast(ref0, 0)
    assert torch.equal(stats, ref0), "running statistics differ between ranks"
    torch.distributed.barrier()
    if rank == 0:
        print("dist_gpu_check ok: world %d, max |synced - concatenated| = %.2e" % (world, worst))
    torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
