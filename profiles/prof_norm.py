"""Driver for ncu: the cfg2 step = fused Fbank (records) + clamp launch + one-launch sentence normalisation."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import speechbrain_b200 as sb
from speechbrain_b200 import ops

torch.manual_seed(0)
B, L = 32, 256000
wav = torch.randn(B, L, device="cuda") * 0.1
lens = torch.linspace(0.6, 1.0, B)[torch.randperm(B)].cuda()
for b in range(B):
    wav[b, int(round(float(lens[b]) * L)):] = 0
tab = sb.Fbank(n_mels=80)._frontend_tables()
for _ in range(3):
    feats, _, ts = ops.fbank_forward(tab, wav, tile_stats=True, top_db=80.0)
    ops.norm_sentence_tiles(feats, lens, ts, out=feats, beside_clamp=True)
    feats, _, ts = ops.fbank_forward(tab, wav, tile_stats=True, top_db=80.0)
    ops.norm_sentence_tiles(feats, lens, ts, out=feats)
    feats, _, ts = ops.fbank_forward(tab, wav, tile_stats=True, top_db=80.0)
    m, s = ops.norm_stats_tiles(feats, lens, ts)
    ops.norm_apply(feats, m, s, per_utt=True, out=feats)
torch.cuda.synchronize()
print("ok")
