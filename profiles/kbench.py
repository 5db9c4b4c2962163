"""Times the fused Fbank kernel alone (cfg2 shape, rotating inputs > L2, CUDA graph of 8
launches, CUDA events).  Usage: python profiles/kbench.py [label]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import speechbrain_b200 as sb
from speechbrain_b200 import ops

label = sys.argv[1] if len(sys.argv) > 1 else ""
torch.manual_seed(0)
pool = [torch.randn(32, 256000, device="cuda") * 0.1 for _ in range(8)]
if os.environ.get("KBENCH_PADDED", "1") == "1":   # bench.py's batch: lengths 0.6-1.0, zero tails
    lens = torch.linspace(0.6, 1.0, 32)[torch.randperm(32)]
    for w in pool:
        for b in range(32):
            w[b, int(round(float(lens[b]) * 256000)):] = 0
fb = sb.Fbank(n_mels=80)
tab = fb._frontend_tables()
for ts in (False, True):
    for w in pool[:2]:
        ops.fbank_forward(tab, w, tile_stats=ts, top_db=None)
    torch.cuda.synchronize()
    keep = []
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for w in pool:
            keep.append(ops.fbank_forward(tab, w, tile_stats=ts, top_db=None))
    for _ in range(3):
        g.replay()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(20):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    # includes the 2 us fill of utt_max per launch
    print("%-28s tile_stats=%d  %.2f us/launch" % (label, ts, e0.elapsed_time(e1) / 160 * 1e3))
