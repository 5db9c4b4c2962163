"""CPU oracle for SpecAugment on features (TEST INFRASTRUCTURE, not product).

numpy restatement of ``speechbrain/lobes/augment.py:32-201`` GIVEN the random draws (the
draws themselves are ``torch.randint`` calls; the product and the tests make them with torch).
The resampling arithmetic lives in PyTorch/ATen (``torch.nn.functional.interpolate`` ->
``upsample_bicubic2d`` / ``upsample_bilinear2d``, align_corners=True; torch>=1.9.0,
reference ``requirements.txt:11``); its published algorithm (aten/src/ATen/native/UpSample.h:
cubic convolution with A = -0.75, border-clamped taps, ``scale = (in-1)/(out-1)``) is
restated here.

Parity status: PINNED on outputs of the unmodified reference class run in the build
container (``tests/golden/specaugment_golden.npz``, generator
``tests/make_golden_augment.py``) - see ``tests/test_oracle_golden.py``.

Only ``tests/`` may import this module.
"""
import numpy as np

_F32 = np.float32


def _cubic1(x, A=-0.75):
    return ((A + 2) * x - (A + 3)) * x * x + 1


def _cubic2(x, A=-0.75):
    return ((A * x - 5 * A) * x + 8 * A) * x - 4 * A


def resample_time(seg, out_n, mode="bicubic", dtype=_F32):
    """seg [B, n, F] -> [B, out_n, F]; F.interpolate(size=(out_n, F), align_corners=True)
    along time (the feature axis keeps its size, so it is the identity there)."""
    seg = np.asarray(seg, dtype=dtype)
    n = seg.shape[1]
    scale = dtype((n - 1) / (out_n - 1)) if out_n > 1 else dtype(0)
    o = np.arange(out_n, dtype=dtype)
    real = (scale * o).astype(dtype)
    if mode == "bicubic":
        fl = np.floor(real)
        i = fl.astype(np.int64)
        t = (real - fl).astype(dtype)
        one = dtype(1)
        w = [_cubic2(t + one), _cubic1(t), _cubic1(one - t), _cubic2((one - t) + one)]
        out = np.zeros((seg.shape[0], out_n, seg.shape[2]), dtype=dtype)
        for k in range(4):
            j = np.clip(i - 1 + k, 0, n - 1)
            out += seg[:, j, :] * w[k].astype(dtype)[None, :, None]
        return out
    if mode == "bilinear":
        i = real.astype(np.int64)
        i1 = i + (i < n - 1)
        l1 = (real - i.astype(dtype)).astype(dtype)
        return (seg[:, i, :] * (dtype(1) - l1)[None, :, None] + seg[:, i1, :] * l1[None, :, None]).astype(dtype)
    raise NotImplementedError(mode)


def time_warp(x, c, w, mode="bicubic", dtype=_F32):
    """lobes/augment.py:116-150 with the drawn centre ``c`` and warped centre ``w``."""
    x = np.asarray(x, dtype=dtype)
    T = x.shape[1]
    out = np.empty_like(x)
    out[:, :w] = resample_time(x[:, :c], w, mode, dtype)
    out[:, w:] = resample_time(x[:, c:], T - w, mode, dtype)
    return out


def mask_along_axis(x, dim, pos, length, fill):
    """lobes/augment.py:152-201 with drawn ``pos`` / ``length`` [B, n_mask]; ``fill`` = 0.0 or
    the mean of x (the caller passes it, :194-196)."""
    x = np.array(x, copy=True)
    D = x.shape[dim]
    ar = np.arange(D)[None, None, :]
    pos = np.asarray(pos)[:, :, None]
    length = np.asarray(length)[:, :, None]
    mask = ((pos <= ar) & (ar < pos + length)).any(axis=1)   # [B, D]
    if dim == 1:
        x[mask] = fill                # rows
    else:
        x[np.broadcast_to(mask[:, None, :], x.shape)] = fill
    return x


def spec_augment(x, warp=None, mode="bicubic", freq=None, time=None, replace_with_zero=True,
                 dtype=_F32):
    """forward (:106-114): warp -> frequency masks -> time masks."""
    x = np.asarray(x, dtype=dtype)
    if warp is not None:
        x = time_warp(x, warp[0], warp[1], mode, dtype)
    if freq is not None:
        fill = 0.0 if replace_with_zero else x.mean(dtype=np.float64).astype(dtype)
        x = mask_along_axis(x, 2, freq[0], freq[1], fill)
    if time is not None:
        fill = 0.0 if replace_with_zero else x.mean(dtype=np.float64).astype(dtype)
        x = mask_along_axis(x, 1, time[0], time[1], fill)
    return x
