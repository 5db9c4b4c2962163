"""Experiment: consecutive (independent) batches alternate between two streams inside one CUDA
graph - does the bandwidth-bound normalisation tail of batch i hide under the issue-bound FFT
kernel of batch i + 1?   python profiles/twostream.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import speechbrain_b200 as sb

torch.manual_seed(0)
B, L = 32, 256000
lens = torch.linspace(0.6, 1.0, B)[torch.randperm(B)].cuda()
pool = []
for i in range(8):
    w = torch.randn(B, L, device="cuda") * 0.1
    for b in range(B):
        w[b, int(round(float(lens[b]) * L)):] = 0
    pool.append(w)
fb, norm = sb.Fbank(n_mels=80), sb.InputNormalization(norm_type="sentence")


from speechbrain_b200 import ops
tab = fb._frontend_tables()
VARIANT = int(os.environ.get("VARIANT", "0"))
REPS = int(os.environ.get("REPS", "1"))      # the pool is walked REPS times inside one graph


def step(w):
    if VARIANT == 0:      # the modules: one-launch normalisation beside the clamp
        return norm(fb(w), lens)
    if VARIANT == 3:      # the FFT kernel alone (records, no clamp launch)
        return ops.fbank_forward(tab, w, tile_stats=True, top_db=None)[0]
    feats, _, ts = ops.fbank_forward(tab, w, tile_stats=True, top_db=80.0)
    if VARIANT == 4:      # + the clamp launch
        return feats
    if VARIANT == 1:      # statistics kernel (one CTA per utterance) + streaming apply: fewer CTA-slots held
        m, sd = ops.norm_stats_tiles(feats, lens, ts)
        return ops.norm_apply(feats, m, sd, per_utt=True)
    m, sd = ops.norm_stats_tiles(feats, lens, ts)
    return ops.norm_apply(feats, m, sd, per_utt=True, out=feats)


def bench(n_streams):
    for w in pool[:2]:
        step(w)
    torch.cuda.synchronize()
    side = [torch.cuda.Stream() for _ in range(n_streams - 1)]
    g, keep = torch.cuda.CUDAGraph(), []
    with torch.cuda.graph(g):
        main = torch.cuda.current_stream()
        ev = torch.cuda.Event()
        ev.record(main)
        for s in side:
            s.wait_event(ev)
        for i, w in enumerate(pool * REPS):
            s = main if i % n_streams == 0 else side[i % n_streams - 1]
            with torch.cuda.stream(s):
                keep.append(step(w))
        for s in side:
            e = torch.cuda.Event()
            e.record(s)
            main.wait_event(e)
    for _ in range(3):
        g.replay()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(25):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / (25 * len(pool) * REPS) * 1e3, keep[:len(pool)]


ref = [step(w).clone() for w in pool]
for n in (1, 2, 3, 4):
    us, keep = bench(n)
    ok = all(torch.equal(torch.nan_to_num(a), torch.nan_to_num(b)) for a, b in zip(keep, ref))
    print("%d stream(s): %6.2f us per step   outputs identical to the single-stream run: %s" % (n, us, ok))
