"""CPU oracle for the acoustic feature frontend (TEST INFRASTRUCTURE, not product).

A numpy restatement of the algorithm that SpeechBrain's ``Fbank`` / ``MFCC`` /
``InputNormalization`` chain computes.  Every function cites the reference
``file:line`` (relative to the reference checkout) whose behaviour it restates.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline /
``--impl reference`` legs may import this package.  The product
(``speechbrain_b200``) never does: it fails loudly when its CUDA library is
missing instead of falling back to this code.

Parity status: PINNED.  ``tests/test_oracle_golden.py`` checks this restatement
against (a) every known-answer vector of the reference's own
``tests/unittests/test_features.py`` and (b) outputs of the unmodified reference
modules run in the build container, committed under ``tests/golden/`` together
with the generating script ``tests/make_golden.py``.

The third-party arithmetic on this path (FFT, GEMM, conv, reductions) lives in
PyTorch/ATen (``torch>=1.9.0``, reference ``requirements.txt:11``); here the
FFT is numpy's pocketfft (native float32 / float64), everything else plain
numpy.  ``dtype`` selects the working precision: ``np.float32`` restates the
reference as it runs; ``np.float64`` is the accuracy anchor (the reference's
fp32-computed constants promoted to double, as SURVEY.md section 8c describes).
"""
import math

import numpy as np

__all__ = [
    "ms_to_samples",
    "hamming_window",
    "num_frames",
    "stft",
    "spectral_magnitude",
    "mel_filter_params",
    "filterbank_matrix",
    "amplitude_to_db",
    "filterbank",
    "dct_matrix",
    "dct",
    "sentence_stats",
    "deltas",
    "context_window",
    "InputNormalizationOracle",
    "fbank",
    "mfcc",
]

_F32 = np.float32


# --------------------------------------------------------------------------
# STFT                                   speechbrain/processing/features.py:50
# --------------------------------------------------------------------------
def ms_to_samples(sample_rate, ms):
    """ms -> samples, ``int(round(sr/1000*ms))``  (processing/features.py:124-129)."""
    return int(round((sample_rate / 1000.0) * ms))


def hamming_window(n, dtype=_F32):
    """Periodic Hamming window, the reference's default ``window_fn``
    (processing/features.py:107,131 -> ``torch.hamming_window(win_length)``):
    ``0.54 - 0.46*cos(2*pi*k/n)`` evaluated in float32 like torch does, then
    promoted to ``dtype``."""
    k = np.arange(n, dtype=_F32)
    w = _F32(0.54) - _F32(0.46) * np.cos(k * _F32(2.0 * math.pi / n), dtype=_F32)
    return w.astype(dtype)


def num_frames(n_samples, hop, n_fft, center=True):
    """Frame count of ``torch.stft`` (processing/features.py:161-172)."""
    padded = n_samples + (2 * (n_fft // 2) if center else 0)
    return 1 + (padded - n_fft) // hop


def stft(
    wav,
    sample_rate=16000,
    win_length=25,
    hop_length=10,
    n_fft=400,
    window=None,
    normalized_stft=False,
    center=True,
    pad_mode="constant",
    onesided=True,
    dtype=_F32,
):
    """``STFT.forward`` (processing/features.py:133-188).

    wav ``[B, L]`` -> ``[B, T, n_fft//2+1, 2]``; ``[B, L, C]`` -> ``[B, T, F, 2, C]``.
    Zero ("constant") centre padding of ``n_fft//2`` per side, window centred
    inside ``n_fft`` when shorter, unnormalised one-sided DFT.
    """
    wav = np.asarray(wav, dtype=dtype)
    multi = wav.ndim == 3
    if multi:  # [B, L, C] -> [B*C, L]            (:143-146)
        b, l, c = wav.shape
        wav = wav.transpose(0, 2, 1).reshape(b * c, l)
    win = ms_to_samples(sample_rate, win_length)
    hop = ms_to_samples(sample_rate, hop_length)
    w = hamming_window(win, dtype) if window is None else np.asarray(window, dtype)
    if win < n_fft:  # torch.stft centres the window inside n_fft
        left = (n_fft - win) // 2
        w = np.pad(w, (left, n_fft - win - left))
    if center:
        mode = {"constant": "constant", "reflect": "reflect",
                "replicate": "edge", "circular": "wrap"}[pad_mode]
        wav = np.pad(wav, ((0, 0), (n_fft // 2, n_fft // 2)), mode=mode)
    t = 1 + (wav.shape[1] - n_fft) // hop
    idx = np.arange(n_fft)[None, :] + hop * np.arange(t)[:, None]
    frames = wav[:, idx] * w  # [B, T, n_fft]
    spec = np.fft.rfft(frames, axis=-1) if onesided else np.fft.fft(frames, axis=-1)
    if normalized_stft:
        spec = spec * dtype(n_fft ** -0.5)
    out = np.stack([spec.real, spec.imag], axis=-1).astype(dtype)  # [B,T,F,2]
    if multi:  # (:175-183) -> [B, T, F, 2, C]
        out = out.reshape(b, c, out.shape[1], out.shape[2], 2).transpose(0, 2, 3, 4, 1)
    return out


def istft(
    spec,
    sample_rate=16000,
    n_fft=None,
    win_length=25,
    hop_length=10,
    window=None,
    center=True,
    sig_length=None,
    dtype=_F32,
):
    """``ISTFT.forward`` (processing/features.py:283-324) = ``torch.istft``: per-frame inverse
    real FFT (1/N), window, overlap-add, division by the overlap-added squared window, centre
    trim of ``n_fft//2`` per side, cut / zero-pad to ``sig_length``.

    spec ``[B, T, F, 2]`` -> ``[B, L]``; ``[B, T, F, 2, C]`` -> ``[B, L, C]`` (:300-307, 320-322).
    ``normalized_stft`` and ``epsilon`` of the reference class never reach ``torch.istft``
    (:309-318), so they have no counterpart here either."""
    spec = np.asarray(spec, dtype=dtype)
    multi = spec.ndim == 5
    if multi:  # [B, T, F, 2, C] -> [B*C, T, F, 2]
        b, t_, f_, _, c = spec.shape
        spec = spec.transpose(0, 4, 1, 2, 3).reshape(b * c, t_, f_, 2)
    if n_fft is None:
        n_fft = (spec.shape[2] - 1) * 2
    win = ms_to_samples(sample_rate, win_length)
    hop = ms_to_samples(sample_rate, hop_length)
    w = hamming_window(win, dtype) if window is None else np.asarray(window, dtype)
    if win < n_fft:
        left = (n_fft - win) // 2
        w = np.pad(w, (left, n_fft - win - left))
    z = spec[..., 0] + 1j * spec[..., 1]
    frames = np.fft.irfft(z, n=n_fft, axis=-1).astype(dtype) * w  # [B, T, n_fft]
    bsz, t = frames.shape[:2]
    full = n_fft + hop * (t - 1)
    y = np.zeros((bsz, full), dtype=dtype)
    env = np.zeros((full,), dtype=dtype)
    for i in range(t):
        y[:, i * hop:i * hop + n_fft] += frames[:, i]
        env[i * hop:i * hop + n_fft] += w * w
    start = n_fft // 2 if center else 0
    if sig_length is not None:
        end = start + sig_length
    else:
        end = full - n_fft // 2 if center else full
    y, env = y[:, start:min(end, full)], env[start:min(end, full)]
    if np.abs(env).min() < 1e-11:
        raise RuntimeError("istft: window overlap-add is zero somewhere (torch.istft raises too)")
    y = y / env
    if end > full:  # requested longer than what the frames cover: zero padding
        y = np.pad(y, ((0, 0), (0, end - full)))
    if multi:
        y = y.reshape(b, c, -1).transpose(0, 2, 1)
    return y.astype(dtype)


def spectral_magnitude(stft_out, power=1, log=False, eps=1e-14):
    """``spectral_magnitude`` (processing/features.py:327-356): sum of squares
    over the LAST axis, ``eps`` added only when ``power < 1``."""
    spectr = (stft_out ** 2).sum(-1)
    if power < 1:
        spectr = spectr + stft_out.dtype.type(eps)
    spectr = spectr ** stft_out.dtype.type(power)
    if log:
        return np.log(spectr + stft_out.dtype.type(eps))
    return spectr


# --------------------------------------------------------------------------
# Filterbank                            speechbrain/processing/features.py:359
# --------------------------------------------------------------------------
def _linspace_f32(start, end, steps):
    """``torch.linspace`` in float32: symmetric evaluation, first half from the
    start, second half from the end (ATen RangeFactories)."""
    start, end = _F32(start), _F32(end)
    if steps == 1:
        return np.array([start], dtype=_F32)
    step = (end - start) / _F32(steps - 1)
    i = np.arange(steps)
    lo = start + step * i.astype(_F32)
    hi = end - step * (steps - 1 - i).astype(_F32)
    return np.where(i < steps // 2, lo, hi).astype(_F32)


def _to_mel(hz):
    """processing/features.py:563-572 (python float math)."""
    return 2595 * math.log10(1 + hz / 700)


def mel_filter_params(n_mels=40, f_min=0, f_max=8000, n_fft=400, sample_rate=16000):
    """Mel-spaced centre frequencies and bands, float32 like the reference
    (processing/features.py:464-488).  Returns ``(f_central[M], band[M],
    all_freqs[n_fft//2+1])``."""
    mel = _linspace_f32(_to_mel(f_min), _to_mel(f_max), n_mels + 2)
    hz = (_F32(700) * (np.power(_F32(10), mel / _F32(2595), dtype=_F32) - _F32(1))).astype(_F32)
    band = (hz[1:] - hz[:-1])[:-1]
    f_central = hz[1:-1]
    all_freqs = _linspace_f32(0, sample_rate // 2, n_fft // 2 + 1)
    return f_central, band, all_freqs


def filterbank_matrix(f_central, band, all_freqs, filter_shape="triangular", dtype=_F32):
    """``[n_stft, n_mels]`` projection matrix.

    triangular  (processing/features.py:586-610): symmetric triangle of
                half-width ``band`` around ``f_central``: max(0, min(s+1, 1-s)),
                s = (f - fc)/band.
    rectangular (:612-635): 1 inside [fc-band, fc+band].
    gaussian    (:637-658): exp(-0.5*((f-fc)/(band/2))**2).
    """
    f = all_freqs.astype(dtype)[:, None]
    fc = f_central.astype(dtype)[None, :]
    bw = band.astype(dtype)[None, :]
    if filter_shape == "triangular":
        slope = (f - fc) / bw
        return np.maximum(dtype(0), np.minimum(slope + dtype(1), dtype(1) - slope))
    if filter_shape == "rectangular":
        return ((f >= fc - bw) & (f <= fc + bw)).astype(dtype)
    return np.exp(dtype(-0.5) * ((f - fc) / (bw / dtype(2))) ** 2)


def amplitude_to_db(x, multiplier=10.0, amin=1e-10, ref_value=1.0, top_db=80.0):
    """``Filterbank._amplitude_to_DB`` (processing/features.py:691-712).  The
    ``top_db`` floor is per leading-axis element: max over the last two axes
    (the whole padded utterance) minus ``top_db``."""
    t = x.dtype.type
    # log10 evaluated in double and rounded once (numpy's float32 log10 is a few
    # ulp off: log10(float32(1e-10)) != -10, which the -100 dB pin needs exact)
    x_db = t(multiplier) * np.log10(np.maximum(x, t(amin)).astype(np.float64)).astype(x.dtype)
    x_db = x_db - t(multiplier * math.log10(max(amin, ref_value)))
    floor = x_db.max(axis=(-2, -1), keepdims=True) - t(top_db)
    if x_db.ndim > 3:  # reference views the floor as [shape[0], 1, 1]
        floor = floor.reshape((x_db.shape[0],) + (1,) * (x_db.ndim - 1))
    return np.maximum(x_db, floor)


def filterbank(
    spec,
    n_mels=40,
    log_mel=True,
    filter_shape="triangular",
    f_min=0,
    f_max=8000,
    n_fft=400,
    sample_rate=16000,
    power_spectrogram=2,
    amin=1e-10,
    ref_value=1.0,
    top_db=80.0,
    mel_params=None,
):
    """``Filterbank.forward`` (processing/features.py:490-560) with frozen,
    un-jittered filters.  ``[B,T,F]`` -> ``[B,T,M]``; ``[B,T,F,C]`` ->
    ``[B,T,M,C]`` (:541-545, 553-558).

    ``mel_params=(f_central, band)`` overrides the numpy-built float32 filter
    constants (the reference's own come from ``torch.linspace`` and ``10**x``,
    whose last-ulp rounding is CPU-kernel dependent and is amplified ~20x by the
    ``10**m - 1`` cancellation at low frequencies)."""
    dtype = spec.dtype.type
    fc, band, freqs = mel_filter_params(n_mels, f_min, f_max, n_fft, sample_rate)
    if mel_params is not None:
        fc, band = (np.asarray(a, dtype=_F32) for a in mel_params)
    mat = filterbank_matrix(fc, band, freqs, filter_shape, dtype)
    multi = spec.ndim == 4
    if multi:
        b, t, f, c = spec.shape
        spec = spec.transpose(0, 3, 1, 2).reshape(b * c, t, f)
    fb = spec @ mat
    if log_mel:
        mult = 10.0 if power_spectrogram == 2 else 20.0  # (:451-454)
        fb = amplitude_to_db(fb, mult, amin, ref_value, top_db)
    if multi:
        fb = fb.reshape(b, c, fb.shape[1], fb.shape[2]).transpose(0, 2, 3, 1)
    return fb


# --------------------------------------------------------------------------
# DCT / Deltas / ContextWindow      speechbrain/processing/features.py:715-937
# --------------------------------------------------------------------------
def dct_matrix(input_size, n_out=20, ortho_norm=True, dtype=_F32):
    """DCT-II matrix ``[input_size, n_out]`` (processing/features.py:740-762),
    built in float32 like the reference, then promoted."""
    if n_out > input_size:
        raise ValueError(
            "Cannot select more DCT coefficients than inputs "
            "(n_out=%i, n_in=%i)" % (n_out, input_size)
        )
    n = np.arange(input_size, dtype=_F32)
    k = np.arange(n_out, dtype=_F32)[:, None]
    m = np.cos(_F32(math.pi / float(input_size)) * (n + _F32(0.5)) * k, dtype=_F32)
    if ortho_norm:
        m[0] *= _F32(1.0 / math.sqrt(2.0))
        m *= _F32(math.sqrt(2.0 / float(input_size)))
    else:
        m *= _F32(2.0)
    return m.T.astype(dtype)


def dct(x, n_out=20, ortho_norm=True):
    """``DCT.forward`` (processing/features.py:764-786), 3-D input."""
    return x @ dct_matrix(x.shape[-1], n_out, ortho_norm, x.dtype.type)


def deltas(x, window_length=5):
    """``Deltas.forward`` (processing/features.py:806-852): regression deltas
    ``sum_k k*x[clamp(t+k)] / (n(n+1)(2n+1)/3)`` with replicate padding in time.
    ``x`` is ``[B, T, F]``."""
    n = (window_length - 1) // 2
    denom = x.dtype.type(n * (n + 1) * (2 * n + 1) / 3)
    t_len = x.shape[1]
    acc = np.zeros_like(x)
    for k in range(-n, n + 1):
        idx = np.clip(np.arange(t_len) + k, 0, t_len - 1)
        acc = acc + x.dtype.type(k) * x[:, idx, :]
    return acc / denom


def context_window(x, left_frames=0, right_frames=0):
    """``ContextWindow.forward`` (processing/features.py:879-937):
    ``out[b, t, f*C + c] = x[b, t + c - left, f]`` with zeros outside
    ``[0, T)``, ``C = left + right + 1``."""
    b, t_len, f = x.shape
    c_len = left_frames + right_frames + 1
    out = np.zeros((b, t_len, f, c_len), dtype=x.dtype)
    for c in range(c_len):
        shift = c - left_frames
        lo, hi = max(0, -shift), min(t_len, t_len - shift)
        if hi > lo:
            out[:, lo:hi, :, c] = x[:, lo + shift:hi + shift, :]
    return out.reshape(b, t_len, f * c_len)


# --------------------------------------------------------------------------
# InputNormalization               speechbrain/processing/features.py:940-1231
# --------------------------------------------------------------------------
def _round_half_even_frames(rel_len, t_len):
    """``torch.round(lengths[i] * T).int()`` (processing/features.py:1023):
    float32 product, round-half-to-even."""
    return int(np.rint(_F32(rel_len) * _F32(t_len)))


def sentence_stats(x, lengths, mean_norm=True, std_norm=True, eps=1e-10):
    """Per-sentence mean / unbiased std over the valid frames, ``[B, F]`` each
    (``_compute_current_stats``, processing/features.py:1120-1145, called per sentence at
    :1020-1031)."""
    means, stds = [], []
    for i in range(x.shape[0]):
        seg = x[i, 0:_round_half_even_frames(lengths[i], x.shape[1])]
        means.append(seg.mean(axis=0) if mean_norm else np.zeros(x.shape[2], x.dtype))
        with np.errstate(invalid="ignore", divide="ignore"):
            std = seg.std(axis=0, ddof=1) if std_norm else np.ones(x.shape[2], x.dtype)
        stds.append(np.maximum(std, x.dtype.type(eps)))
    return np.stack(means), np.stack(stds)


class InputNormalizationOracle:
    """Stateful restatement of ``InputNormalization`` (processing/features.py:941).

    ``forward`` mirrors :1000-1118, ``_stats`` mirrors :1120-1145.  ``sentence``
    and ``speaker`` modes write into ``x`` and return it; ``batch`` / ``global``
    return a new array.
    """

    def __init__(self, mean_norm=True, std_norm=True, norm_type="global",
                 avg_factor=None, update_until_epoch=3):
        self.mean_norm, self.std_norm = mean_norm, std_norm
        self.norm_type, self.avg_factor = norm_type, avg_factor
        self.update_until_epoch = update_until_epoch
        self.glob_mean = self.glob_std = 0  # torch.tensor([0]) in the reference
        self.spk_mean, self.spk_std, self.spk_count = {}, {}, {}
        self.count, self.weight, self.eps = 0, 1.0, 1e-10
        self.training = True

    def _stats(self, seg):
        t = seg.dtype.type
        if self.mean_norm:
            mean = seg.mean(axis=0)
        else:
            mean = np.zeros(1, seg.dtype)
        if self.std_norm:
            with np.errstate(invalid="ignore", divide="ignore"):
                std = seg.std(axis=0, ddof=1)  # torch.std is unbiased
        else:
            std = np.ones(1, seg.dtype)
        return mean, np.maximum(std, t(self.eps))  # NaN propagates like torch.max

    def forward(self, x, lengths, spk_ids=None, epoch=0):
        n_b, t_len = x.shape[0], x.shape[1]
        means, stds = [], []
        for i in range(n_b):
            n_valid = _round_half_even_frames(lengths[i], t_len)
            mean, std = self._stats(x[i, 0:n_valid])
            means.append(mean)
            stds.append(std)
            if self.norm_type == "sentence":
                x[i] = (x[i] - mean) / std
            elif self.norm_type == "speaker":
                spk = int(np.asarray(spk_ids[i]).reshape(-1)[0])
                if self.training:
                    if spk not in self.spk_mean:
                        self.spk_mean[spk], self.spk_std[spk] = mean, std
                        self.spk_count[spk] = 1
                    else:
                        self.spk_count[spk] += 1
                        w = (1 / self.spk_count[spk]) if self.avg_factor is None else self.avg_factor
                        self.weight = w
                        self.spk_mean[spk] = (1 - w) * self.spk_mean[spk] + w * mean
                        self.spk_std[spk] = (1 - w) * self.spk_std[spk] + w * std
                    s_mean, s_std = self.spk_mean[spk], self.spk_std[spk]
                elif spk in self.spk_mean:
                    s_mean, s_std = self.spk_mean[spk], self.spk_std[spk]
                else:
                    s_mean, s_std = mean, std
                x[i] = (x[i] - s_mean.astype(x.dtype)) / s_std.astype(x.dtype)
        if self.norm_type in ("batch", "global"):
            cur_mean = np.mean(np.stack(means), axis=0)
            cur_std = np.mean(np.stack(stds), axis=0)
            if self.norm_type == "batch":
                return (x - cur_mean) / cur_std
            if self.training:
                if self.count == 0:
                    self.glob_mean, self.glob_std = cur_mean, cur_std
                elif epoch < self.update_until_epoch:
                    w = (1 / (self.count + 1)) if self.avg_factor is None else self.avg_factor
                    self.weight = w
                    t = x.dtype.type
                    self.glob_mean = t(1 - w) * self.glob_mean + t(w) * cur_mean
                    self.glob_std = t(1 - w) * self.glob_std + t(w) * cur_std
                self.count += 1
            with np.errstate(divide="ignore", invalid="ignore"):
                return ((x - self.glob_mean) / self.glob_std).astype(x.dtype)
        return x

    __call__ = forward


# --------------------------------------------------------------------------
# Lobes                                     speechbrain/lobes/features.py:22,150
# --------------------------------------------------------------------------
def fbank(wav, deltas_=False, context=False, sample_rate=16000, f_min=0, f_max=None,
          n_fft=400, n_mels=40, filter_shape="triangular", left_frames=5,
          right_frames=5, win_length=25, hop_length=10, dtype=_F32, window=None,
          mel_params=None):
    """``Fbank.forward`` (lobes/features.py:130-147)."""
    if f_max is None:
        f_max = sample_rate / 2  # (:105-106)
    s = stft(wav, sample_rate, win_length, hop_length, n_fft, window=window, dtype=dtype)
    fb = filterbank(spectral_magnitude(s), n_mels=n_mels, filter_shape=filter_shape,
                    f_min=f_min, f_max=f_max, n_fft=n_fft, sample_rate=sample_rate,
                    mel_params=mel_params)
    if deltas_:
        d1 = deltas(fb)
        fb = np.concatenate([fb, d1, deltas(d1)], axis=2)
    if context:
        fb = context_window(fb, left_frames, right_frames)
    return fb


def mfcc(wav, deltas_=True, context=True, sample_rate=16000, f_min=0, f_max=None,
         n_fft=400, n_mels=23, n_mfcc=20, filter_shape="triangular", left_frames=5,
         right_frames=5, win_length=25, hop_length=10, dtype=_F32, window=None,
         mel_params=None):
    """``MFCC.forward`` (lobes/features.py:263-281)."""
    if f_max is None:
        f_max = sample_rate / 2
    s = stft(wav, sample_rate, win_length, hop_length, n_fft, window=window, dtype=dtype)
    fb = filterbank(spectral_magnitude(s), n_mels=n_mels, filter_shape=filter_shape,
                    f_min=f_min, f_max=f_max, n_fft=n_fft, sample_rate=sample_rate,
                    mel_params=mel_params)
    out = dct(fb, n_mfcc)
    if deltas_:
        d1 = deltas(out)
        out = np.concatenate([out, d1, deltas(d1)], axis=2)
    if context:
        out = context_window(out, left_frames, right_frames)
    return out
