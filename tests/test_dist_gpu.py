"""Multi-rank GPU test of the path's one collective (BASELINE.json configs[4]):
two ranks each run Fbank(80) + InputNormalization("global", sync_stats=True) on THEIR (uneven)
shard for three training steps - the real kernels (fused Fbank with tile records + clamp launch, norm_stats_tiles,
norm_reduce, norm_update, norm_apply) and a real all-reduce of the 161-float payload: NCCL with
one rank per GPU when the box has two GPUs, else both ranks on cuda:0 over gloo.  Every rank's
output must equal its slice of the ORACLE (numpy, float64 features) run single-process on the
concatenated global batch, and the running statistics must be identical on all ranks."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import dist_workers as dw  # noqa: E402
from oracle import frontend_np as onp  # noqa: E402

pytestmark = pytest.mark.gpu


def test_global_norm_all_reduce_equals_oracle_on_concatenated_batch():
    assert torch.cuda.is_available()
    world = 2
    backend = "nccl" if torch.cuda.device_count() >= world else "gloo"
    port = 29700 + (os.getpid() % 2000)
    out = mp.Manager().dict()
    mp.spawn(dw.gpu_worker, args=(world, port, out, backend), nprocs=world, join=True)
    import speechbrain_b200 as sb

    fb = sb.Fbank(n_mels=80)
    consts = (fb.compute_fbanks.f_central.numpy(), fb.compute_fbanks.band.numpy())
    ref = onp.InputNormalizationOracle(norm_type="global")
    for step in range(dw.STEPS):
        ws, ls = zip(*[dw.wav_shard(r, step) for r in range(world)])
        feats = onp.fbank(np.concatenate(ws), n_mels=80, dtype=np.float64, mel_params=consts,
                          window=fb.compute_STFT.window.numpy())
        want = ref(feats, np.concatenate(ls), epoch=0)
        lo = 0
        for r in range(world):
            got = out[r][0][step]
            err = np.abs(got - want[lo:lo + got.shape[0]]).max()
            assert err < 1e-3, (backend, step, r, err)   # dB error of the fp32 FFT over std ~ 5
            lo += got.shape[0]
    for r in range(world):
        np.testing.assert_allclose(out[r][1], ref.glob_mean, rtol=1e-5, atol=2e-4)
        np.testing.assert_allclose(out[r][2], ref.glob_std, rtol=1e-4, atol=2e-4)
        assert out[r][3] == dw.STEPS and out[r][4], "running statistics differ between ranks"
    print("dist gpu test ok: backend %s" % backend)
