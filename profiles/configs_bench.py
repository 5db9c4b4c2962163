"""Device-resident timing of the five BASELINE.json configs on one GPU (per-GPU shard for
cfg5), CUDA graphs over rotating inputs, with the algorithmic-byte roofline fraction.

    python profiles/configs_bench.py [--json out.json]
Under torchrun (N ranks) cfg5 runs with sync_stats=True (NCCL all-reduce of the payload)."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import speechbrain_b200 as sb

PEAK = 6553.3
p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
if os.path.exists(p):
    PEAK = float(json.load(open(p))["hbm_gbs"])

rank = int(os.environ.get("RANK", 0))
world = int(os.environ.get("WORLD_SIZE", 1))
local = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    torch.distributed.init_process_group("nccl", device_id=dev)


def cfgs():
    yield ("cfg1 Fbank-40 8x1s", 8, 16000, 40, lambda: (sb.Fbank(), None))
    yield ("cfg2 Fbank-80+sentence norm 32x16s", 32, 256000, 80,
           lambda: (sb.Fbank(n_mels=80), sb.InputNormalization(norm_type="sentence")))
    yield ("cfg3 MFCC-660 64x4s", 64, 64000, 660, lambda: (sb.MFCC(), None))
    yield ("cfg4 Fbank-80+sentence mean norm 256x3s", 256, 48000, 80,
           lambda: (sb.Fbank(n_mels=80), sb.InputNormalization(norm_type="sentence", std_norm=False)))
    yield ("cfg5 Fbank-80+global norm 128x30s per GPU", 128, 480000, 80,
           lambda: (sb.Fbank(n_mels=80),
                    sb.InputNormalization(norm_type="global", sync_stats=world > 1).train()))


rows = []
for name, B, L, F, make in cfgs():
    torch.manual_seed(1234 + rank)
    fe, norm = make()
    if norm is not None and norm.norm_type == "sentence":
        pass   # unmodified recipe flow: nothing to declare (round 1 paired the modules here)
    T = 1 + L // 160
    n_in = max(2, min(8, int(300e6 // (4 * B * L)) + 1))  # rotate over > L2 worth of input
    pool = [torch.randn(B, L, device=dev) * 0.1 for _ in range(n_in)]
    lens = torch.linspace(0.6, 1.0, B, device=dev)

    def step(w):
        x = fe(w)
        return norm(x, lens, epoch=0) if norm is not None else x

    for w in pool[:2]:
        step(w)
    torch.cuda.synchronize()
    keep = []
    g = torch.cuda.CUDAGraph()
    try:
        with torch.cuda.graph(g):
            for w in pool:
                keep.append(step(w))
        run = g.replay
    except Exception as e:  # NCCL capture unavailable: run eagerly
        run = lambda: [step(w) for w in pool]  # noqa: E731
    for _ in range(3):
        run()
    reps = max(3, int(200 // n_in))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        run()
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / (reps * n_in) * 1e3
    # independent batches alternating between two streams inside the graph (bench.py's default);
    # not for the global statistics of cfg5, whose running update orders the steps
    us2 = None
    if norm is None or norm.norm_type == "sentence":
        side = torch.cuda.Stream()
        g2, keep2 = torch.cuda.CUDAGraph(), []
        with torch.cuda.graph(g2):
            main = torch.cuda.current_stream()
            fork = torch.cuda.Event()
            fork.record(main)
            side.wait_event(fork)
            for i, w in enumerate(pool):
                with torch.cuda.stream(main if i % 2 == 0 else side):
                    keep2.append(step(w))
            join = torch.cuda.Event()
            join.record(side)
            main.wait_event(join)
        for _ in range(3):
            g2.replay()
        torch.cuda.synchronize()
        e0.record()
        for _ in range(reps):
            g2.replay()
        e1.record()
        torch.cuda.synchronize()
        us2 = e0.elapsed_time(e1) / (reps * n_in) * 1e3
        del g2, keep2
    alg = 4 * B * L + 4 * B * T * F
    rows.append({"config": name, "us_per_step": us, "audio_s_per_s": B * L / 16000 / (us * 1e-6),
                 "alg_bytes": alg, "gbs": alg / us / 1e3, "hbm_frac": alg / us / 1e3 / PEAK,
                 "us_per_step_two_streams": us2,
                 "hbm_frac_two_streams": None if us2 is None else alg / us2 / 1e3 / PEAK, "n_gpus": world})
    del pool, keep, g
    torch.cuda.empty_cache()
if rank == 0:
    for r in rows:
        two = "" if r["us_per_step_two_streams"] is None else "   two streams: %6.1f us = %4.1f%%" % (
            r["us_per_step_two_streams"], 100 * r["hbm_frac_two_streams"])
        print("%-44s %9.1f us  %12.0f audio-s/s  %7.1f GB/s  %5.1f%% of %.0f GB/s%s" % (
            r["config"], r["us_per_step"], r["audio_s_per_s"], r["gbs"], 100 * r["hbm_frac"], PEAK, two))
    if "--json" in sys.argv:
        json.dump(rows, open(sys.argv[sys.argv.index("--json") + 1], "w"), indent=1)
if world > 1:
    torch.distributed.destroy_process_group()
