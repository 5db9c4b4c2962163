import sys, os
sys.path.insert(0, "/root/repo")
import torch, speechbrain_b200 as sb
which = sys.argv[1]
torch.manual_seed(0)
if which == "cfg3":
    wav = torch.randn(64, 64000, device="cuda") * 0.1
    fe = sb.MFCC()
    for _ in range(3): out = fe(wav)
else:
    wav = torch.randn(128, 480000, device="cuda") * 0.1
    lens = torch.linspace(0.6, 1.0, 128, device="cuda")
    fe = sb.Fbank(n_mels=80); norm = sb.InputNormalization(norm_type="global").train()
    for _ in range(3): out = norm(fe(wav), lens, epoch=0)
torch.cuda.synchronize(); print("ok")
