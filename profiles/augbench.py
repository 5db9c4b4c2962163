"""SpecAugment on the cfg2 feature tensor [32, 1601, 80]: this repo's kernel (ops level and
module level) and the torch-op chain of the reference run on the same GPU, CUDA events.
    python profiles/augbench.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import speechbrain_b200 as sb
from speechbrain_b200 import ops
from test_specaugment import _reference_on_device

B, T, F = 32, 1601, 80
pool = [torch.randn(B, T, F, device="cuda") for _ in range(12)]   # 12 x 16.4 MB > L2
aug = sb.SpecAugment()


def timed(fn, reps=10, graph=False):
    for x in pool[:3]:
        fn(x)
    if graph:   # kernel-level numbers: no host launch overhead in the timed region
        torch.cuda.synchronize()
        g, keep = torch.cuda.CUDAGraph(), []
        with torch.cuda.graph(g):
            for x in pool:
                keep.append(fn(x))
        g.replay()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(); e0.record()
        for _ in range(reps):
            g.replay()
        e1.record(); torch.cuda.synchronize()
        return e0.elapsed_time(e1) / (reps * len(pool)) * 1e3
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(reps):
        for x in pool:
            fn(x)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / (reps * len(pool)) * 1e3


fm = (torch.randint(0, 60, (B, 2), device="cuda"), torch.randint(0, 20, (B, 2), device="cuda"))
tm = (torch.randint(0, 1500, (B, 2), device="cuda"), torch.randint(0, 100, (B, 2), device="cuda"))
alg = 2 * 4 * B * T * F
us = timed(lambda x: ops.specaugment_forward(x, (800, 803), "bicubic", fm, tm, 0.0), graph=True)
print("kernel warp+masks (out of place)   %7.2f us  %6.0f GB/s (read + write once)" % (us, alg / us / 1e3))
us = timed(lambda x: ops.specaugment_forward(x, None, "bicubic", fm, tm, 0.0), graph=True)
print("kernel masks only (in place)       %7.2f us" % us)
us = timed(lambda x: aug(x))
print("module SpecAugment() incl. draws   %7.2f us" % us)
us = timed(lambda x: _reference_on_device(x, aug))
print("reference torch-op chain, same GPU %7.2f us" % us)
