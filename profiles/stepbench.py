"""Cumulative timing of the cfg2 step's kernels (CUDA graphs over 8 rotating inputs):
fused kernel without / with records + in-launch clamp | + statistics from the tile records | the step.
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
    if upto == 1:     # mode (b) without a clamp: what the records alone cost
        return ops.fbank_forward(tab, w, tile_stats=True, top_db=None)[0]
    feats, _, ts = ops.fbank_forward(tab, w, tile_stats=True, top_db=80.0)   # + the clamp launch
    if upto == 3:
        ops.norm_stats_tiles(feats, lens, ts)
    if upto == 4:
        ops.norm_sentence_tiles(feats, lens, ts, out=feats)
    if upto == 7:     # truly in place: the held tiles go through the ticket (csrc/norm.cu)
        ops.norm_sentence_tiles(feats, lens, ts, out=feats, beside_clamp=True)
    if upto == 8:     # what InputNormalization.forward does: a fresh tensor, x takes over its storage
        return ops.norm_sentence_tiles(feats, lens, ts, beside_clamp=True)
    if upto == 5:
        mean, std = ops.norm_stats_tiles(feats, lens, ts)
        ops.norm_apply(feats, mean, std, per_utt=True, out=feats)
    if upto == 6:     # the statistics the generic way (no records): 2 launches + apply
        mean, std = ops.norm_stats(feats, lens)
        ops.norm_apply(feats, mean, std, per_utt=True, out=feats)
    return feats


prev = 0.0
for upto, name in enumerate(["fbank400_v3 mode (a): no records, no clamp (+memset)",
                             "fbank400_v3 mode (b): records, no clamp",
                             "fbank400_v3 mode (b): records + clamp launch",
                             "  + norm_stats_tiles", "  + norm_sentence_tiles, in place",
                             "  + norm_stats_tiles + norm_apply",
                             "  + generic stats + apply",
                             "  + norm_sentence_tiles beside the clamp, in place",
                             "  + norm_sentence_tiles beside the clamp, new tensor (THE STEP)"]):
    base = {3: 2, 4: 2, 5: 2, 6: 2, 7: 2, 8: 2}.get(upto)
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
    res = globals().setdefault("res", {})
    res[upto] = us
    print("%-52s %7.2f us/step%s" % (name, us, "  (%+.2f)" % (us - res[base]) if base is not None else ""))
