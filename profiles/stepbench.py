"""Cumulative timing of the cfg2 step's kernels (CUDA graphs over 8 rotating inputs):
fused kernel without / with its finisher | + statistics from the prefix records | + apply.
python profiles/stepbench.py"""
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
    if upto == 0:     # mode (a): un-clamped output, utterance maxima via atomics (needs a memset)
        return ops.fbank_forward(tab, w)[0]
    feats, _, ts = ops.fbank_forward(tab, w, tile_stats=True, top_db=80.0)   # finisher mode
    if upto == 2:
        ops.norm_stats_prefix(feats, lens, ts)
    if upto == 3:
        ops.norm_sentence_prefix(feats, lens, ts, out=feats)
    if upto == 4:     # the statistics the generic way (no records): 2 launches + apply
        mean, std = ops.norm_stats(feats, lens)
        ops.norm_apply(feats, mean, std, per_utt=True, out=feats)
    return feats


prev = 0.0
for upto, name in enumerate(["fbank400_v3, no finisher (+memset)", "fbank400_v3 finisher (clamp+prefix)",
                             "+ norm_stats_prefix", "+ stats_prefix + norm_apply (the step)",
                             "finisher + generic stats + apply"]):
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
    print("%-42s %7.2f us/step  (%+.2f)" % (name, us, us - prev))
    prev = us
