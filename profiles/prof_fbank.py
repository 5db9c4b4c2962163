"""Small driver for ncu: a few launches of the cfg2 frontend step (32 x 16 s)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import speechbrain_b200 as sb

torch.manual_seed(0)
wav = torch.randn(32, 256000, device="cuda") * 0.1
lens = torch.linspace(0.6, 1.0, 32, device="cuda")
if os.environ.get("PROF_PADDED", "1") == "1":   # bench.py's batch: zero tails; 0: lengths = 1
    for b in range(32):
        wav[b, int(round(float(lens[b]) * 256000)):] = 0
fb, norm = sb.Fbank(n_mels=80), sb.InputNormalization(norm_type="sentence")
for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 and sys.argv[1].isdigit() else 4):
    out = norm(fb(wav), lens)
torch.cuda.synchronize()
print("ok", float(out.abs().mean()))
