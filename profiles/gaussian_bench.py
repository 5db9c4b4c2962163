import sys, os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import speechbrain_b200 as sb
wav = torch.randn(32, 256000).cuda() * 0.1
def bench(fb, tag):
    for _ in range(3): y = fb(wav)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): y = fb(wav)
    e1.record(); torch.cuda.synchronize()
    print("%-44s %8.1f us per call" % (tag, e0.elapsed_time(e1) / 20 * 1e3)); return y
for M in (80, 40):
    os.environ["SBFE_MEL_MMA"] = "0"
    a = bench(sb.Fbank(n_mels=M, filter_shape="gaussian"), "gaussian-%d n_fft 400, sparse kernels" % M)
    os.environ["SBFE_MEL_MMA"] = "2"
    b = bench(sb.Fbank(n_mels=M, filter_shape="gaussian"), "gaussian-%d n_fft 400, STFT + tensor-core mel" % M)
    fbt = sb.Fbank(n_mels=M, filter_shape="gaussian"); fbt.compute_fbanks.mel_precision = "tf32"
    c = bench(fbt, "gaussian-%d n_fft 400, STFT + TF32 mel" % M)
    print("   max |3xTF32 - sparse| %.2e   max |TF32 - sparse| %.2e" % (float((a-b).abs().max()), float((a-c).abs().max())))
    os.environ["SBFE_MEL_MMA"] = "0"
    a = bench(sb.Fbank(n_mels=M, n_fft=512, filter_shape="gaussian"), "gaussian-%d n_fft 512, sparse kernels" % M)
    os.environ["SBFE_MEL_MMA"] = "2"
    b = bench(sb.Fbank(n_mels=M, n_fft=512, filter_shape="gaussian"), "gaussian-%d n_fft 512, STFT + tensor-core mel" % M)
    print("   max |3xTF32 - sparse| %.2e" % float((a-b).abs().max()))
os.environ["SBFE_MEL_MMA"] = "2"
b = bench(sb.Fbank(n_mels=80), "triangular-80 n_fft 400, STFT + tensor-core mel")
del os.environ["SBFE_MEL_MMA"]
a = bench(sb.Fbank(n_mels=80), "triangular-80 n_fft 400, fused kernel")
print("   max diff %.2e" % float((a-b).abs().max()))
