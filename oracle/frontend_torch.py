"""CPU baseline port of the reference frontend on torch CPU ops (TEST / BENCH
INFRASTRUCTURE, not product).

The reference executes this path as a chain of ATen calls (torch.stft ->
pow/sum -> matmul -> log10/clamp/amax/max -> per-sentence mean/std loop); this
file restates that chain op for op so that ``bench.py``'s ``cpu_baseline`` and
``--impl reference`` legs time what the reference's own CPU path costs on the
GPU box's host cores (the reference package itself cannot travel there).
Each function cites the reference lines it follows.  Parity: pinned against the
committed reference outputs in tests/test_oracle_golden.py (same fixtures as
the numpy oracle).  Never imported by ``speechbrain_b200``.
"""
import math

import torch


def _tables(n_mels, n_fft, sample_rate, f_min, f_max):
    # Filterbank.__init__, processing/features.py:464-488
    to_mel = lambda hz: 2595 * math.log10(1 + hz / 700)  # noqa: E731
    mel = torch.linspace(to_mel(f_min), to_mel(f_max), n_mels + 2)
    hz = 700 * (10 ** (mel / 2595) - 1)
    band = (hz[1:] - hz[:-1])[:-1]
    f_central = hz[1:-1]
    n_stft = n_fft // 2 + 1
    freqs = torch.linspace(0, sample_rate // 2, n_stft).repeat(n_mels, 1)
    return f_central, band, freqs


def fbank(wav, n_mels=40, n_fft=400, sample_rate=16000, win_ms=25, hop_ms=10, f_min=0,
          f_max=None, top_db=80.0, amin=1e-10):
    """Fbank.forward with default (triangular, log-mel) filters,
    lobes/features.py:130-147."""
    if f_max is None:
        f_max = sample_rate / 2
    win = int(round(sample_rate / 1000.0 * win_ms))
    hop = int(round(sample_rate / 1000.0 * hop_ms))
    window = torch.hamming_window(win)
    # STFT.forward, processing/features.py:161-186
    st = torch.view_as_real(torch.stft(wav, n_fft, hop, win, window, True, "constant", False, True,
                                       return_complex=True)).transpose(2, 1)
    # spectral_magnitude, :347-356
    power = st.pow(2).sum(-1)
    # Filterbank.forward: the matrix is rebuilt on every call, :499-536, 586-610
    f_central, band, freqs = _tables(n_mels, n_fft, sample_rate, f_min, f_max)
    fc = f_central.repeat(freqs.shape[1], 1).transpose(0, 1)
    bw = band.repeat(freqs.shape[1], 1).transpose(0, 1)
    slope = (freqs - fc) / bw
    mat = torch.max(torch.zeros(1), torch.min(slope + 1.0, -slope + 1.0)).transpose(0, 1)
    fb = torch.matmul(power, mat)
    # _amplitude_to_DB, :691-712
    x_db = 10 * torch.log10(torch.clamp(fb, min=amin))
    x_db -= 10 * math.log10(max(amin, 1.0))
    floor = x_db.amax(dim=(-2, -1)) - top_db
    return torch.max(x_db, floor.view(x_db.shape[0], 1, 1))


def sentence_norm(x, lengths, mean_norm=True, std_norm=True, eps=1e-10):
    """InputNormalization(norm_type='sentence').forward, the per-sentence Python
    loop of processing/features.py:1020-1035 and :1120-1145 (in place)."""
    for i in range(x.shape[0]):
        n = torch.round(lengths[i] * x.shape[1]).int()
        seg = x[i, 0:n, ...]
        mean = torch.mean(seg, dim=0) if mean_norm else torch.tensor([0.0])
        std = torch.std(seg, dim=0) if std_norm else torch.tensor([1.0])
        std = torch.max(std, eps * torch.ones_like(std))
        x[i] = (x[i] - mean) / std
    return x


def fbank_sentence_norm(wav, lengths, n_mels=80):
    """BASELINE cfg2: Fbank(n_mels=80) + InputNormalization('sentence')."""
    return sentence_norm(fbank(wav, n_mels=n_mels), lengths)
