"""Host time (python + ctypes + launches, no synchronisation) of the module calls, eager."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import speechbrain_b200 as sb

torch.manual_seed(0)
B, L = 32, 256000
w = torch.randn(B, L, device="cuda") * 0.1
lens = torch.linspace(0.6, 1.0, B, device="cuda")
fb, norm = sb.Fbank(n_mels=80), sb.InputNormalization(norm_type="sentence")
for _ in range(5):
    norm(fb(w), lens)
torch.cuda.synchronize()
N = 200
t_fb = t_norm = 0.0
for _ in range(N):
    t0 = time.perf_counter()
    f = fb(w)
    t1 = time.perf_counter()
    y = norm(f, lens)
    t2 = time.perf_counter()
    t_fb += t1 - t0
    t_norm += t2 - t1
    if _ % 20 == 19:
        torch.cuda.synchronize()
print("host time per call: Fbank %.1f us, InputNormalization %.1f us" % (t_fb / N * 1e6, t_norm / N * 1e6))
import cProfile, pstats
pr = cProfile.Profile()
pr.enable()
for _ in range(100):
    y = norm(fb(w), lens)
pr.disable()
torch.cuda.synchronize()
pstats.Stats(pr).sort_stats("tottime").print_stats(22)
