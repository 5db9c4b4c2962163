"""Cumulative timing of the cfg2 step's kernels (CUDA graphs over 8 rotating inputs):
fbank | +clamp_tiles | +norm_stats_tiles | +norm_apply.  python profiles/stepbench.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import speechbrain_b200 as sb
from speechbrain_b200 import ops

torch.manual_seed(0)
B, L = 32, 256000
pool = [torch.randn(B, L, device="cuda") * 0.1 for _ in range(8)]
lens = torch.linspace(0.6, 1.0, B)[torch.randperm(B)].cuda()
for i, w in enumerate(pool):
    for b in range(B):
        w[b, int(round(float(lens[b]) * L)):] = 0
fb = sb.Fbank(n_mels=80)
tab = fb._frontend_tables()


def step(w, upto):
    feats, _, ts = ops.fbank_forward(tab, w, tile_stats=True)
    if 1 <= upto <= 3:
        ops.clamp_tiles(feats, ts, 80.0)
    if upto == 2:
        mean, std = ops.norm_stats_tiles(feats, lens, ts)
    if upto == 3:
        mean, std = ops.norm_stats_tiles(feats, lens, ts)
        ops.norm_apply(feats, mean, std, per_utt=True, out=feats, stats_mark=ops.stats_launch_mark(feats))
    if upto == 4:   # pair(): no clamp launch, the normalisation clamps on the fly
        ts.top_db = 80.0
        mean, std, keys = ops.norm_stats_tiles(feats, lens, ts, want_max=True)
        ops.norm_apply(feats, mean, std, per_utt=True, out=feats, utt_max=keys, top_db=80.0,
                       stats_mark=ops.stats_launch_mark(feats))
    return feats


prev = 0.0
for upto, name in enumerate(["fbank400_v3 (+tile records)", "+ clamp_tiles", "+ norm_stats_tiles", "+ norm_apply",
                             "paired: fbank + stats + apply(clamp)"]):
    for w in pool[:2]:
        step(w, upto)
    torch.cuda.synchronize()
    keep = []
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for w in pool:
            keep.append(step(w, upto))
    for _ in range(3):
        g.replay()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(20):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / 160 * 1e3
    print("%-30s %7.2f us/step  (+%.2f)" % (name, us, us - prev))
    prev = us
