"""CPU restatement (numpy) of the time-domain augmentation data movements - TEST INFRASTRUCTURE,
never imported by ``speechbrain_b200``.

Given the draws (chunks, filter taps, clip value, polyphase tables), what the reference modules do
to the samples:
  drop_chunks   DropChunk.forward          speechbrain/processing/speech_augmentation.py:1155-1171
  fir_same      convolve1d(padding = pad)  speechbrain/processing/signal_processing.py:157-277
                (as DropFreq.forward applies it, speech_augmentation.py:1003-1007)
  clip          DoClip.forward             speech_augmentation.py:1222-1226
  resample      Resample._perform_resample speech_augmentation.py:600-665
Pinned on outputs of the unmodified reference modules (tests/golden/timedomain_golden.npz,
tests/make_golden_timedomain.py) in tests/test_timedomain.py.
"""
import numpy as np


def drop_chunks(x, rounds, noise=None, noise_max=None):
    """x [B, L]; rounds[j] = [(utterance, start, length, noise offset), ...] applied round by round
    (python slice semantics: clipped at the ends)."""
    out = np.array(x, copy=True)
    for chunks in rounds:
        for b, s, n, off in chunks:
            if noise is None:
                out[b, s:s + n] = 0.0
            else:
                seg = out[b, s:s + n]
                nm = noise_max[b]
                seg[:] = (2 * nm * noise[off:off + len(seg)] - nm).astype(out.dtype)
    return out


def fir_same(x, taps, pad, dtype=np.float64):
    """out[b][t] = sum_k taps[k] x[b][t + k - pad], x zero outside (conv1d: no kernel flip)."""
    x = np.asarray(x, dtype=dtype)
    taps = np.asarray(taps, dtype=dtype)
    B, L = x.shape
    K = len(taps)
    xp = np.zeros((B, L + K - 1), dtype=dtype)
    xp[:, pad:pad + L] = x
    out = np.zeros((B, L), dtype=dtype)
    for k in range(K):
        out += taps[k] * xp[:, k:k + L]
    return out


def clip(x, value):
    return np.clip(x, -value, value)


def resample(x, weights, first, in_stride, n_out, dtype=np.float64):
    """Output sample n * P + i = sum_k weights[i][k] x[first[i] + n * in_stride + k], x zero outside."""
    x = np.asarray(x, dtype=dtype)
    w = np.asarray(weights, dtype=dtype)
    B, L = x.shape
    P, W = w.shape
    out = np.zeros((B, n_out), dtype=dtype)
    for i in range(P):
        o = np.arange(i, n_out, P)
        if len(o) == 0:
            continue
        base = int(first[i]) + (o // P) * in_stride
        for k in range(W):
            g = base + k
            ok = (g >= 0) & (g < L)
            out[:, o[ok]] += w[i, k] * x[:, g[ok]]
    return out
