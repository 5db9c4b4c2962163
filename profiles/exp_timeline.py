"""Timeline of the cfg2 step's three kernels from %globaltimer stamps (experiment build
-DSBFE_EXP_TIMELINE: first CTA start / first wait-return / last CTA end per kernel).
Run through profiles/exp_timeline.sh."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import speechbrain_b200 as sb
from speechbrain_b200 import _lib, ops

torch.manual_seed(0)
B, L = 32, 256000
pool = [torch.randn(B, L, device="cuda") * 0.1 for _ in range(4)]
lens = torch.linspace(0.6, 1.0, B)[torch.randperm(B)].cuda()
for w in pool:
    for b in range(B):
        w[b, int(round(float(lens[b]) * L)):] = 0
tab = sb.Fbank(n_mels=80)._frontend_tables()
h = _lib.lib()
h.sbfe_exp_timeline.argtypes = [C.c_void_p, C.c_int]
buf = (C.c_ulonglong * 24)()
names = ["producer", "clamp", "norm", "norm: statistics done"]
for beside in (True, False):
    def step(w):
        f, _, ts = ops.fbank_forward(tab, w, tile_stats=True, top_db=80.0)
        return ops.norm_sentence_tiles(f, lens, ts, beside_clamp=beside)   # a new tensor, like the module
    for w in pool:
        step(w)
    keep = []
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        keep.append(step(pool[0]))          # ONE step per graph: the stamps are per kernel, not per step
    for _ in range(3):
        g.replay()
    rows = []
    for rep in range(20):
        h.sbfe_exp_timeline(buf, 1)
        g.replay()
        torch.cuda.synchronize()
        h.sbfe_exp_timeline(buf, 0)
        v = list(buf)
        t0 = v[1]                            # producer's wait-return = start of the step's work
        rows.append([(v[3 * k + i] - t0) / 1e3 for k in range(8) for i in range(3)])
    med = [sorted(r[i] for r in rows)[len(rows) // 2] for i in range(24)]
    print("beside_clamp=%s   (us after the producer's wait returned; median of 20 single-step graphs)" % beside)
    for k in range(3):
        print("  %-10s first CTA in %7.2f   first wait-return %7.2f   last CTA out %7.2f"
              % (names[k], med[3 * k], med[3 * k + 1], med[3 * k + 2]))
    print("  %-10s first CTA: floor known %7.2f   per-thread merges done %7.2f   statistics done %7.2f"
          % ("norm", med[12], med[15], med[9]))
