"""CPU restatement (numpy) of the time-domain augmentation data movements - TEST INFRASTRUCTURE,
never imported by ``speechbrain_b200``.

Given the draws (chunks, filter taps, clip value, polyphase tables), what the reference modules do
to the samples:
  drop_chunks   DropChunk.forward          speechbrain/processing/speech_augmentation.py:1155-1171
  fir_same      convolve1d(padding = pad)  speechbrain/processing/signal_processing.py:157-277
                (as DropFreq.forward applies it, speech_augmentation.py:1003-1007)
  clip          DoClip.forward             speech_augmentation.py:1222-1226
  resample      Resample._perform_resample speech_augmentation.py:600-665
  add_noise     AddNoise.forward (white)   speech_augmentation.py:126-168 (+ compute_amplitude,
                dB_to_amplitude: signal_processing.py:51-56, 355-372)
  add_babble    AddBabble.forward          speech_augmentation.py:861-906
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


def add_noise(x, noise, lengths, snr_db, normalize=False):
    """x, noise [B, L] float32; lengths [B] relative; snr_db [B].  The reference's float32 ops in
    its order: amp = sum|x| / (lengths * L); f = 1 / (10^(snr/20) + 1); y = x * (1 - f);
    y += (f * amp) * noise; normalize: y / max(max|y|, 1)."""
    f32 = np.float32
    x = np.asarray(x, dtype=f32)
    noise = np.asarray(noise, dtype=f32)
    B, L = x.shape
    ln = (np.asarray(lengths, dtype=f32) * f32(L)).reshape(B, 1)
    amp = np.abs(x).sum(axis=1, keepdims=True, dtype=np.float64).astype(f32) / ln
    snr = np.asarray(snr_db, dtype=f32).reshape(B, 1)
    f = (f32(1) / (np.power(f32(10), snr / f32(20), dtype=f32) + f32(1))).astype(f32)
    y = x * (f32(1) - f)
    y = y + (f * amp) * noise
    if normalize:
        y = y / np.maximum(np.abs(y).max(axis=1, keepdims=True), f32(1))
    return y.astype(f32)


def add_babble(x, lengths, snr_db, speakers):
    """x [B, L] float32; lengths [B] relative; snr_db [B].  The reference's float32 ops in its
    order: babble = roll(x, 1) + roll(x, 2) + ... (in place, in that order); babble_len = the
    running maximum of the rolled lengths; scale = (f amp) / (sum|babble| / babble_len + 1e-14)."""
    f32 = np.float32
    x = np.asarray(x, dtype=f32)
    B, L = x.shape
    ln = (np.asarray(lengths, dtype=f32) * f32(L)).reshape(B, 1)
    amp = np.abs(x).sum(axis=1, keepdims=True, dtype=np.float64).astype(f32) / ln
    snr = np.asarray(snr_db, dtype=f32).reshape(B, 1)
    f = (f32(1) / (np.power(f32(10), snr / f32(20), dtype=f32) + f32(1))).astype(f32)
    y = x * (f32(1) - f)
    babble = np.roll(x, 1, axis=0).copy()
    blen = np.roll(ln, 1, axis=0)
    for i in range(1, speakers):
        babble = babble + np.roll(x, 1 + i, axis=0)
        blen = np.maximum(blen, np.roll(blen, 1, axis=0))
    bamp = np.abs(babble).sum(axis=1, keepdims=True, dtype=np.float64).astype(f32) / blen
    babble = babble * ((f * amp) / (bamp + f32(1e-14)))
    return (y + babble).astype(f32)
