"""Tensor-level wrappers over the C ABI (include/sbfe.h).

Each function takes / returns torch tensors, launches on
``torch.cuda.current_stream()`` and never synchronises.  torch is plumbing only
(device memory + streams); all arithmetic is in libsbfe.so.  CPU tensors raise:
there is no fallback path.
"""
import ctypes as C

import torch

from . import _lib
from ._lib import FbankDesc, check, lib

__all__ = [
    "FrontendTables", "fbank_forward", "stft_forward", "istft_forward", "spectral_magnitude",
    "filterbank_forward", "post_forward", "norm_stats", "norm_stats_tiles", "norm_sentence_tiles",
    "TileStats", "norm_reduce", "norm_update", "norm_apply", "decode_utt_max", "mark_written", "version_of",
]


def _require_cuda(t, name):
    if not isinstance(t, torch.Tensor):
        raise TypeError("%s must be a torch.Tensor" % name)
    if not t.is_cuda:
        raise RuntimeError(
            "%s is on %s: speechbrain_b200 runs on CUDA only (no CPU fallback; keep the "
            "reference modules for CPU)" % (name, t.device))
    if t.requires_grad and torch.is_grad_enabled():
        raise RuntimeError(
            "%s requires grad: the B200 frontend is forward-only (SURVEY.md N4)" % name)


def _f32c(t, name):
    """float32 + contiguous; half/bfloat16 are promoted like the reference's
    window multiply does, float64 is rejected like its matmul does."""
    _require_cuda(t, name)
    if t.dtype == torch.float64:
        raise TypeError("%s is float64; the frontend computes in float32" % name)
    if t.dtype != torch.float32:
        t = t.float()
    return t.contiguous()


_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)


def _stream():
    """The current stream of the current device as a raw pointer (one C call where available:
    the eager step is host-bound, torch.cuda.current_stream() costs ~4 us per use)."""
    if _raw_stream is not None:
        return C.c_void_p(_raw_stream(torch.cuda.current_device()))
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def stream_id(device):
    """The raw handle (int) of ``device``'s current stream."""
    with _on_device(device):
        return _stream().value or 0


class _on_device:
    """`with _on_device(d)` only when d is not already current (the common case costs
    ~8 us of context management per call otherwise)."""
    __slots__ = ("ctx",)

    def __init__(self, device):
        idx = device.index if device.index is not None else torch.cuda.current_device()
        self.ctx = None if idx == torch.cuda.current_device() else torch.cuda.device(idx)

    def __enter__(self):
        if self.ctx is not None:
            self.ctx.__enter__()

    def __exit__(self, *a):
        if self.ctx is not None:
            self.ctx.__exit__(*a)


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


class FrontendTables:
    """Host-built constant tables of one STFT(+filterbank) configuration, moved
    to each device on first use (the reference re-copies them on every call:
    processing/features.py:166, 534-536)."""

    def __init__(self, n_fft, hop, window, fbank_matrix=None, log_mel=True, multiplier=10.0,
                 amin=1e-10, db_multiplier=0.0, mel_precision=0):
        self.mel_precision = int(mel_precision)   # sbfe.h: 0 = 3xTF32 (fp32 grade), 1 = TF32
        self.n_fft, self.hop = int(n_fft), int(hop)
        self.bins = self.n_fft // 2 + 1
        win = window.detach().to("cpu", torch.float32)
        if win.numel() > self.n_fft:
            raise ValueError("win_length (%d samples) > n_fft (%d)" % (win.numel(), n_fft))
        if win.numel() < self.n_fft:  # torch.stft centres a short window inside n_fft
            left = (self.n_fft - win.numel()) // 2
            win = torch.nn.functional.pad(win, (left, self.n_fft - win.numel() - left))
        self.h_window = win.contiguous()
        self.log_mel = bool(log_mel)
        self.multiplier, self.amin = float(multiplier), float(amin)
        self.db_offset = float(multiplier) * float(db_multiplier)
        # what the reference's float32 ops give for x == amin (features.py:701-702)
        a = torch.tensor(self.amin, dtype=torch.float32)
        fl = self.multiplier * torch.log10(torch.clamp(a, min=self.amin))
        fl -= self.db_offset
        self.floor_db = float(fl)
        self.n_mels = 0
        self._h_mel = None
        if fbank_matrix is not None:
            self._set_matrix(fbank_matrix)
        self._dev = {}

    def _set_matrix(self, mat):
        mat = mat.detach().to("cpu", torch.float32)
        if mat.shape[0] != self.bins:
            raise ValueError("filterbank matrix has %d rows, expected %d" % (mat.shape[0], self.bins))
        self.n_mels = int(mat.shape[1])
        # pair-uniform padded layout (sbfe_fbank_desc.mel_pad4 == 2): filters (2k, 2k+1) store
        # the same number G of 4-weight groups, zero filled, every offset a multiple of 4, and
        # start + 4 G never leaves the zero-extended spectrum (rows = 4 * ceil(bins / 4)) - a
        # filter that would is moved down and gets leading zeros instead
        rows = 4 * ((self.bins + 3) // 4)
        spans = []
        for m in range(self.n_mels):
            nz = torch.nonzero(mat[:, m]).flatten()
            spans.append((0, 0) if nz.numel() == 0 else (int(nz[0]), int(nz[-1]) - int(nz[0]) + 1))
        start, count, offset, weights, fill = [], [], [], [], 0
        for m in range(self.n_mels):
            mate = m ^ 1
            groups = max((spans[m][1] + 3) // 4,
                         (spans[mate][1] + 3) // 4 if mate < self.n_mels else 0)
            s, c = spans[m]
            shift = max(0, s + 4 * groups - rows)
            s, c = s - shift, c + shift
            start.append(s)
            count.append(c)
            offset.append(fill)
            weights.append(torch.cat([mat[s:s + c, m], torch.zeros(4 * groups - c)]))
            fill += 4 * groups
        w = torch.cat(weights) if weights and fill else torch.zeros(4)
        self._h_mel = (torch.tensor(start, dtype=torch.int32), torch.tensor(count, dtype=torch.int32),
                       torch.tensor(offset, dtype=torch.int32), w.contiguous())
        self.nnz = int(fill)

    def on(self, device):
        key = (device.type, device.index)
        hit = self._dev.get(key)
        if hit is not None:
            return hit
        d_window = self.h_window.to(device)
        keep = [d_window]
        desc = FbankDesc()
        desc.n_fft, desc.hop, desc.n_mels = self.n_fft, self.hop, self.n_mels
        desc.log_mel = int(self.log_mel)
        desc.multiplier, desc.amin = self.multiplier, self.amin
        desc.db_offset, desc.floor_db = self.db_offset, self.floor_db
        desc.h_window, desc.d_window = self.h_window.data_ptr(), d_window.data_ptr()
        if self._h_mel is not None:
            dev = [t.to(device) for t in self._h_mel]
            keep += dev
            desc.d_mel_start, desc.d_mel_count = dev[0].data_ptr(), dev[1].data_ptr()
            desc.d_mel_offset, desc.d_mel_weight = dev[2].data_ptr(), dev[3].data_ptr()
            desc.mel_nnz = self.nnz
            desc.mel_pad4 = 2
            desc.mel_precision = self.mel_precision
        entry = (desc, keep)
        self._dev[key] = entry
        return entry


def _wav2d(wav):
    wav = _f32c(wav, "wav")
    if wav.dim() != 2:
        raise ValueError("expected a [batch, time] waveform, got shape %s" % (tuple(wav.shape),))
    return wav


def mark_written(t):
    """Tell torch that ``t`` was written in place by a kernel launched through ctypes (which
    autograd's version counter cannot see): bumps ``t._version``, so that anything keyed on it -
    the one-shot tile records handed from Fbank to InputNormalization - sees the change."""
    torch.autograd.graph.increment_version(t)
    return t


def version_of(t):
    """``t._version``, or None where torch keeps none (inference tensors)."""
    try:
        return t._version
    except RuntimeError:
        return None


class TileStats:
    """Tile records emitted by the fused Fbank kernel (include/sbfe.h: sbfe_fbank_forward (b)):
    record j = per-column {mean, M2} of the final (clamped) features over the existing frames of
    tile j.  ``version`` = ``_version`` of the feature tensor they describe; they are void once
    that tensor is written to."""

    __slots__ = ("stats", "frames", "version", "top_db", "stream")

    def __init__(self, stats, frames, top_db=None):
        self.stats, self.frames, self.version, self.top_db = stats, frames, None, top_db
        self.stream = None   # the stream the producing call went into


def ragged_supported(tables, device):
    """True when sbfe_fbank_forward_ragged can take this configuration."""
    desc, _ = tables.on(device)
    return int(lib().sbfe_fbank_ragged_supported(C.byref(desc))) > 0


def fbank_forward(tables, wav, tile_stats=False, top_db=80.0):
    """wav [B, L] (or a dataio.RaggedBatch) -> mel features [B, T, M].

    tile_stats=False: (dB or linear mel WITHOUT the top_db clamp, utt_max keys [B] int32).
    tile_stats=True (where the configuration has tile records): (FINAL features - clamped
    to ``top_db`` below each utterance's maximum (None: no clamp) -, None, TileStats); where it
    has none: (un-clamped, utt_max, None).  sbfe_fbank_forward[_ragged[_pcm16]]."""
    ragged = None
    if hasattr(wav, "packed"):
        ragged = wav
        _require_cuda(ragged.packed, "wav")
        if ragged.packed.dtype not in (torch.float32, torch.int16) or not ragged_supported(
                tables, ragged.device):
            wav, ragged = ragged.to_padded(), None
    if ragged is not None:
        B, L, dev_ = ragged.batch_size, ragged.max_len, ragged.device
        wav = ragged.packed
    else:
        wav = _wav2d(wav)
        B, L = wav.shape
    T = 1 + L // tables.hop
    desc, _ = tables.on(wav.device)
    out = torch.empty((B, T, tables.n_mels), dtype=torch.float32, device=wav.device)
    L_ = lib()
    info = tables.__dict__.setdefault("_call_info", {})     # per (device, B, L): library answers
    key = (wav.device.index, B, L)
    hit = info.get(key)
    if hit is None:
        if len(info) > 64:
            info.clear()
        fr = int(L_.sbfe_fbank_tile_frames(C.byref(desc)))
        hit = info[key] = (int(L_.sbfe_fbank_workspace_bytes(C.byref(desc), B, L)), fr,
                           int(L_.sbfe_fbank_tile_stats_floats(C.byref(desc))) if fr > 0 else 0)
    ws_bytes, frames_, words = hit
    ws = torch.empty((ws_bytes,), dtype=torch.uint8, device=wav.device) if ws_bytes > 0 else None
    ts, frames = None, 0
    if tile_stats and B > 0:
        frames = frames_
        if frames > 0:
            ts = torch.empty((B, (T + frames - 1) // frames, words), dtype=torch.float32,
                             device=wav.device)
    # with tile records the utterance maximum travels in the records: no atomics, no memset
    utt_max = None if ts is not None else torch.zeros((B,), dtype=torch.int32, device=wav.device)
    db = -1.0 if top_db is None else float(top_db)
    with _on_device(wav.device):
        if ragged is not None and wav.dtype == torch.int16:
            check(L_.sbfe_fbank_forward_ragged_pcm16(C.byref(desc), _ptr(wav), _ptr(ragged.offsets),
                                                     _ptr(ragged.nsamples), B, L, ragged.scale,
                                                     _ptr(out), _ptr(utt_max), _ptr(ts), db, _stream()))
        elif ragged is not None:
            check(L_.sbfe_fbank_forward_ragged(C.byref(desc), _ptr(wav), _ptr(ragged.offsets),
                                               _ptr(ragged.nsamples), B, L, _ptr(out),
                                               _ptr(utt_max), _ptr(ts), db, _stream()))
        else:
            check(L_.sbfe_fbank_forward(C.byref(desc), _ptr(wav), B, L, L, _ptr(out),
                                        _ptr(utt_max), _ptr(ts), db, _ptr(ws),
                                        max(ws_bytes, 0), _stream()))
    if tile_stats:
        rec = None
        if ts is not None:
            rec = TileStats(ts, frames, top_db)
            rec.stream = stream_id(wav.device)
        return out, utt_max, rec
    return out, utt_max


def _stats_out(x, lengths):
    B, T, F = x.shape
    lengths = lengths.to(device=x.device, dtype=torch.float32).contiguous()
    mean = torch.empty((B, F), dtype=torch.float32, device=x.device)
    std = torch.empty((B, F), dtype=torch.float32, device=x.device)
    return lengths, mean, std


def norm_stats_tiles(x, lengths, ts, mean_norm=True, std_norm=True, eps=1e-10):
    """norm_stats from the fused kernel's tile records (2 F floats per tile + the rows of the
    tile the valid length cuts; no pass over x).  sbfe_norm_stats_tiles."""
    B, T, F = x.shape
    lengths, mean, std = _stats_out(x, lengths)
    with _on_device(x.device):
        check(lib().sbfe_norm_stats_tiles(_ptr(x), _ptr(lengths), B, T, F, int(bool(mean_norm)),
                                          int(bool(std_norm)), float(eps), _ptr(ts.stats), ts.frames,
                                          -1.0 if ts.top_db is None else float(ts.top_db),
                                          _ptr(mean), _ptr(std), _stream()))
    return mean, std


def norm_sentence_tiles(x, lengths, ts, mean_norm=True, std_norm=True, eps=1e-10, out=None,
                        beside_clamp=False, out_dtype=None):
    """Sentence normalisation straight from the tile records: statistics + (x - mean) / std in
    ONE launch (sbfe_norm_sentence_tiles).  In place when ``out is x``.  ``beside_clamp``: nothing
    that writes x, the records or lengths was launched since the Fbank call that made them
    (sbfe.h: SBFE_BESIDE_CLAMP) - the kernel then runs beside that call's clamp launch."""
    B, T, F = x.shape
    lengths = lengths.to(device=x.device, dtype=torch.float32).contiguous()
    bf16 = out_dtype == torch.bfloat16
    if out is None or bf16:   # a bfloat16 result (sbfe.h: SBFE_OUT_BF16) is always a new tensor
        out = torch.empty_like(x, dtype=torch.bfloat16 if bf16 else None)
    with _on_device(x.device):
        check(lib().sbfe_norm_sentence_tiles(_ptr(x), _ptr(lengths), B, T, F, int(bool(mean_norm)),
                                             int(bool(std_norm)), float(eps), _ptr(ts.stats),
                                             ts.frames, -1.0 if ts.top_db is None else float(ts.top_db),
                                             (1 if beside_clamp else 0) | (2 if bf16 else 0),
                                             _ptr(out), _stream()))
    return mark_written(out)


def stft_forward(tables, wav, power_only=False, center=True, reflect=False, onesided=True):
    """wav [B, L] -> [B, T, bins, 2] (or [B, T, bins] power).  sbfe_stft_forward; the keyword
    flags are torch.stft's center / pad_mode="reflect" / onesided."""
    wav = _wav2d(wav)
    B, L = wav.shape
    flags = (0 if center else 1) | (2 if reflect and center else 0) | (0 if onesided else 4)
    desc, _ = tables.on(wav.device)
    T = int(lib().sbfe_stft_frames(C.byref(desc), L, flags))
    if T <= 0 and not center:
        raise RuntimeError("STFT(center=False) needs at least n_fft = %d samples, got %d"
                           % (tables.n_fft, L))
    bins = tables.bins if onesided else tables.n_fft
    shape = (B, T, bins) if power_only else (B, T, bins, 2)
    out = torch.empty(shape, dtype=torch.float32, device=wav.device)
    with _on_device(wav.device):
        check(lib().sbfe_stft_forward(C.byref(desc), _ptr(wav), B, L, L, _ptr(out),
                                      int(power_only), flags, _stream()))
    return out


def istft_forward(spec, n_fft, hop, window, center, out_len):
    """spec [B, T, n_fft//2+1, 2] -> [B, out_len] (sbfe_istft_forward).  ``window``: device
    float32 [n_fft]."""
    spec = _f32c(spec, "stft")
    if spec.dim() != 4 or spec.shape[-1] != 2 or spec.shape[2] != n_fft // 2 + 1:
        raise ValueError("expected [batch, frames, n_fft//2+1, 2], got %s" % (tuple(spec.shape),))
    B, T = spec.shape[:2]
    out = torch.empty((B, out_len), dtype=torch.float32, device=spec.device)
    L_ = lib()
    ws = torch.empty((max(int(L_.sbfe_istft_workspace_bytes(B, T, n_fft)), 16),), dtype=torch.uint8,
                     device=spec.device)
    with _on_device(spec.device):
        check(L_.sbfe_istft_forward(_ptr(spec), B, T, int(n_fft), int(hop), _ptr(window),
                                    int(bool(center)), int(out_len), _ptr(out), _ptr(ws), ws.numel(),
                                    _stream()))
    return out


def spectral_magnitude(stft, power=1, log=False, eps=1e-14):
    stft = _f32c(stft, "stft")
    if stft.shape[-1] != 2:
        raise ValueError("last dimension must hold (re, im)")
    out = torch.empty(stft.shape[:-1], dtype=torch.float32, device=stft.device)
    with _on_device(stft.device):
        check(lib().sbfe_spectral_magnitude(_ptr(stft), out.numel(), float(power), int(bool(log)),
                                            float(eps), _ptr(out), _stream()))
    return out


def filterbank_forward(tables, spec):
    """spec [B, T, bins] -> (mel [B, T, M] without the clamp, utt_max keys)."""
    spec = _f32c(spec, "spectrogram")
    B, T, F = spec.shape
    if F != tables.bins:
        raise ValueError("spectrogram has %d bins, filterbank expects %d" % (F, tables.bins))
    desc, _ = tables.on(spec.device)
    out = torch.empty((B, T, tables.n_mels), dtype=torch.float32, device=spec.device)
    utt_max = torch.zeros((B,), dtype=torch.int32, device=spec.device)
    with _on_device(spec.device):
        check(lib().sbfe_filterbank_forward(C.byref(desc), _ptr(spec), B, T, _ptr(out),
                                            _ptr(utt_max), _stream()))
    return out, utt_max


_DENSE_TABLES = {}


def filterbank_forward_dense(spec, mat, log_mel=True, multiplier=10.0, amin=1e-10, db_multiplier=0.0):
    """Filterbank projection with a dense DEVICE matrix ``mat`` [bins, M] (a learnable filterbank
    rebuilds it every step: no host copy, no sparsity scan): spec [B, T, bins] -> (dB or linear
    mel [B, T, M] without the top_db clamp, utt_max keys).  Same entry point as filterbank_forward
    with a descriptor whose tables describe the full matrix, which routes to the tensor-core
    kernel (sbfe.h: mel_precision)."""
    _require_cuda(spec, "spectrogram")
    spec = _f32c(spec, "spectrogram")
    B, T, bins = spec.shape
    M = mat.shape[1]
    if mat.shape[0] != bins:
        raise ValueError("filterbank matrix has %d rows, spectrogram has %d bins" % (mat.shape[0], bins))
    wt = mat.t().contiguous().float()                      # [M][bins]: filter m = row m
    key = (spec.device, bins, M)
    idx = _DENSE_TABLES.get(key)
    if idx is None:
        idx = (torch.zeros(M, dtype=torch.int32, device=spec.device),
               torch.full((M,), bins, dtype=torch.int32, device=spec.device),
               (torch.arange(M, dtype=torch.int32, device=spec.device) * bins).contiguous())
        _DENSE_TABLES[key] = idx
    tab = FrontendTables(2 * (bins - 1), 1, torch.ones(2 * (bins - 1)), None, log_mel, multiplier,
                         amin, db_multiplier)
    desc = FbankDesc()
    desc.n_fft, desc.hop, desc.n_mels, desc.log_mel = tab.n_fft, 1, M, int(bool(log_mel))
    desc.multiplier, desc.amin = tab.multiplier, tab.amin
    desc.db_offset, desc.floor_db = tab.db_offset, tab.floor_db
    desc.h_window = desc.d_window = wt.data_ptr()          # not used by the projection
    desc.d_mel_start, desc.d_mel_count = idx[0].data_ptr(), idx[1].data_ptr()
    desc.d_mel_offset, desc.d_mel_weight = idx[2].data_ptr(), wt.data_ptr()
    desc.mel_nnz, desc.mel_pad4, desc.mel_precision = M * bins, 0, 0
    out = torch.empty((B, T, M), dtype=torch.float32, device=spec.device)
    utt_max = torch.zeros((B,), dtype=torch.int32, device=spec.device)
    with _on_device(spec.device):
        check(lib().sbfe_filterbank_forward(C.byref(desc), _ptr(spec), B, T, _ptr(out), _ptr(utt_max),
                                            _stream()))
    return out, utt_max


def post_forward(x, utt_max=None, top_db=80.0, dct=None, deltas=False, delta_window=5,
                 context=False, left=0, right=0, inplace=False):
    """clamp -> DCT -> deltas -> context window (sbfe_post_forward)."""
    x = _f32c(x, "x")
    B, T, F = x.shape
    pure_clamp = dct is None and not deltas and not (context and (left or right))
    F0 = dct.shape[1] if dct is not None else F
    F_out = F0 * (3 if deltas == 1 or deltas is True else 1) * ((left + right + 1) if context else 1)
    out = x if (inplace and pure_clamp) else torch.empty((B, T, F_out), dtype=torch.float32,
                                                         device=x.device)
    with _on_device(x.device):
        check(lib().sbfe_post_forward(_ptr(x), B, T, F, _ptr(utt_max), float(top_db), _ptr(dct),
                                      int(F0 if dct is not None else 0), int(deltas),
                                      int(delta_window), int(bool(context)), int(left), int(right),
                                      _ptr(out), _stream()))
    return mark_written(out) if out is x else out


def norm_stats(x, lengths, mean_norm=True, std_norm=True, eps=1e-10, utt_max=None, top_db=80.0):
    """Per-utterance mean / unbiased std over the valid frames -> ([B,F], [B,F])."""
    B, T, F = x.shape
    lengths = lengths.to(device=x.device, dtype=torch.float32).contiguous()
    mean = torch.empty((B, F), dtype=torch.float32, device=x.device)
    std = torch.empty((B, F), dtype=torch.float32, device=x.device)
    L_ = lib()
    ws_bytes = L_.sbfe_norm_workspace_bytes(B, T, F)
    ws = torch.empty((max(ws_bytes, 16),), dtype=torch.uint8, device=x.device)
    with _on_device(x.device):
        check(L_.sbfe_norm_stats(_ptr(x), _ptr(lengths), B, T, F, int(bool(mean_norm)),
                                 int(bool(std_norm)), float(eps), _ptr(utt_max), float(top_db),
                                 _ptr(mean), _ptr(std), _ptr(ws), ws_bytes, _stream()))
    return mean, std


def norm_reduce(mean, std):
    """[B,F] stats -> [2F+1] payload {sum mean, sum std, B} (all-reduce-able)."""
    B, F = mean.shape
    cur = torch.empty((2 * F + 1,), dtype=torch.float32, device=mean.device)
    with _on_device(mean.device):
        check(lib().sbfe_norm_reduce(_ptr(mean), _ptr(std), B, F, _ptr(cur), _stream()))
    return cur


def norm_update(cur, glob_mean, glob_std, first, weight):
    F = glob_mean.numel()
    with _on_device(cur.device):
        check(lib().sbfe_norm_update(_ptr(cur), F, int(bool(first)), float(weight),
                                     _ptr(glob_mean), _ptr(glob_std), _stream()))


def norm_speaker_update(mean, std, row, mode, weight, table_mean, table_std):
    """Speaker-mode running statistics (sbfe_norm_speaker_update): returns the [B, F] rows to
    normalise with; updates the table rows in place."""
    B, F = mean.shape
    out_m, out_s = torch.empty_like(mean), torch.empty_like(std)
    with _on_device(mean.device):
        check(lib().sbfe_norm_speaker_update(_ptr(mean), _ptr(std), B, F, _ptr(row), _ptr(mode),
                                             _ptr(weight), _ptr(table_mean), _ptr(table_std),
                                             _ptr(out_m), _ptr(out_s), _stream()))
    return out_m, out_s


def norm_apply(x, mean, std, per_utt, out=None, utt_max=None, top_db=80.0, out_dtype=None):
    """(max(x, floor) - mean) / std (sbfe_norm_apply); in place when ``out is x``;
    ``out_dtype=torch.bfloat16``: a new bfloat16 tensor (sbfe.h: SBFE_OUT_BF16)."""
    B, T, F = x.shape
    bf16 = out_dtype == torch.bfloat16
    if out is None or bf16:
        out = torch.empty_like(x, dtype=torch.bfloat16 if bf16 else None)
    with _on_device(x.device):
        check(lib().sbfe_norm_apply(_ptr(x), B, T, F, _ptr(mean), _ptr(std), int(bool(per_utt)),
                                    _ptr(utt_max), float(top_db), 2 if bf16 else 0, _ptr(out),
                                    _stream()))
    return mark_written(out) if out is x else out


def decode_utt_max(keys):
    """int32 keys -> float32 maxima on the host (tests / debugging only)."""
    import struct

    out = []
    for k in (keys.to(torch.int64) & 0xFFFFFFFF).tolist():
        bits = (k & 0x7FFFFFFF) if (k & 0x80000000) else (~k & 0xFFFFFFFF)
        out.append(struct.unpack("<f", struct.pack("<I", bits))[0])
    return torch.tensor(out, dtype=torch.float32)


def tensor_mean(x):
    """x.mean() as a device scalar [1] (sbfe_mean; no host round trip)."""
    x = _f32c(x, "x")
    L_ = lib()
    ws = torch.empty((int(L_.sbfe_mean_workspace_bytes()),), dtype=torch.uint8, device=x.device)
    out = torch.empty((1,), dtype=torch.float32, device=x.device)
    with _on_device(x.device):
        check(L_.sbfe_mean(_ptr(x), x.numel(), _ptr(out), _ptr(ws), ws.numel(), _stream()))
    return out


def specaugment_forward(x, warp=None, warp_mode="bicubic", freq_masks=None, time_masks=None,
                        fill=0.0):
    """SpecAugment data movement (sbfe_specaugment_forward).  x [B, T, F] float32 contiguous.

    warp = (c, w) python ints or None; freq_masks / time_masks = (pos, len) int64 [B, n] device
    tensors or None; fill = python float or a device scalar tensor.  With a warp a new tensor is
    returned; without one the masked elements of ``x`` are overwritten and ``x`` is returned."""
    _require_cuda(x, "x")
    if x.dtype != torch.float32 or not x.is_contiguous() or x.dim() != 3:
        raise TypeError("specaugment_forward needs a contiguous float32 [B, T, F] tensor")
    B, T, F = x.shape
    modes = {"bicubic": 0, "bilinear": 1}
    if warp is not None and warp_mode not in modes:
        raise NotImplementedError("time_warp_mode=%r (bicubic and bilinear are built)" % (warp_mode,))

    def pair(m):
        if m is None:
            return None, None, 0
        pos, ln = m
        pos = pos.reshape(B, -1).to(device=x.device, dtype=torch.int64).contiguous()
        ln = ln.reshape(B, -1).to(device=x.device, dtype=torch.int64).contiguous()
        return pos, ln, pos.shape[1]

    fpos, flen, nf = pair(freq_masks)
    tpos, tlen, nt = pair(time_masks)
    d_fill, f_fill = (fill, 0.0) if isinstance(fill, torch.Tensor) else (None, float(fill))
    out = torch.empty_like(x) if warp is not None else x
    c, w = (int(warp[0]), int(warp[1])) if warp is not None else (-1, -1)
    with _on_device(x.device):
        check(lib().sbfe_specaugment_forward(_ptr(x), B, T, F, c, w, modes.get(warp_mode, 0),
                                             _ptr(fpos), _ptr(flen), nf, _ptr(tpos), _ptr(tlen),
                                             nt, _ptr(d_fill), f_fill, _ptr(out), _stream()))
    return mark_written(out) if out is x else out


# ---- time-domain augmentation (csrc/timedomain.cu; processing/speech_augmentation.py) ----------
def wav_drop_chunks(out, chunks, noise=None, noise_max=None):
    """Zero (or noise-fill) ``chunks`` = [(utterance, start, length, noise offset), ...] of
    ``out`` [B, L] in place (sbfe_wav_drop_chunks); chunks of one call must not overlap inside an
    utterance."""
    B, L = out.shape
    tab = torch.tensor(chunks, dtype=torch.int32).reshape(-1, 4)
    max_len = int(tab[:, 2].max()) if tab.numel() else 0
    d_tab = tab.to(out.device, non_blocking=True)
    with _on_device(out.device):
        check(lib().sbfe_wav_drop_chunks(_ptr(out), B, L, 1, _ptr(d_tab), tab.shape[0], max_len,
                                         _ptr(noise), _ptr(noise_max), _stream()))
    return mark_written(out)


def wav_fir(x, taps, pad):
    """out[b][t] = sum_k taps[k] x[b][t + k - pad] with zero padding (sbfe_wav_fir)."""
    B, L = x.shape
    out = torch.empty_like(x)
    taps = taps.to(device=x.device, dtype=torch.float32).contiguous()
    with _on_device(x.device):
        check(lib().sbfe_wav_fir(_ptr(x), B, L, _ptr(taps), taps.numel(), int(pad), _ptr(out),
                                 _stream()))
    return out


def wav_notch_taps(freqs, width, filter_len, device):
    """DropFreq's composite band-reject taps [filter_len] on ``device`` from the drawn
    frequencies (a short python list; sbfe_wav_notch_taps: they travel as launch arguments)."""
    taps = torch.empty(filter_len, dtype=torch.float32, device=device)
    arr = (C.c_float * max(len(freqs), 1))(*freqs)
    with _on_device(device):
        check(lib().sbfe_wav_notch_taps(arr, len(freqs), float(width), int(filter_len), _ptr(taps),
                                        _stream()))
    return taps


def wav_clip(x, value):
    """clamp(x, -value, value) into a new tensor (sbfe_wav_clip)."""
    out = torch.empty_like(x)
    with _on_device(x.device):
        check(lib().sbfe_wav_clip(_ptr(x), x.numel(), float(value), _ptr(out), _stream()))
    return out


def wav_add_noise(x, noise, lengths, snr_db, normalize=False):
    """AddNoise's arithmetic for a drawn noise tensor (sbfe_wav_abs_sums + sbfe_wav_add_noise
    [+ sbfe_wav_peak_normalize]): x, noise [B, L], lengths (relative) and snr_db [B] on the device;
    returns a new tensor."""
    B, L = x.shape
    out = torch.empty_like(x)
    if B == 0 or L == 0:
        return out
    L_ = lib()
    chunks = int(L_.sbfe_wav_abs_chunks(L))
    partials = torch.empty((B, chunks), dtype=torch.float32, device=x.device)
    absmax = torch.zeros(B, dtype=torch.int32, device=x.device) if normalize else None
    with _on_device(x.device):
        check(L_.sbfe_wav_abs_sums(_ptr(x), B, L, _ptr(partials), _stream()))
        check(L_.sbfe_wav_add_noise(_ptr(x), _ptr(noise), B, L, _ptr(partials), _ptr(lengths),
                                    _ptr(snr_db), _ptr(out), _ptr(absmax) if normalize else None,
                                    _stream()))
        if normalize:
            check(L_.sbfe_wav_peak_normalize(_ptr(out), B, L, _ptr(absmax), _stream()))
    return out


def wav_add_babble(x, lengths, snr_db, speakers):
    """AddBabble's arithmetic (sbfe_wav_add_babble): x [B, L], lengths (relative) and snr_db [B] on
    the device; returns a new tensor."""
    B, L = x.shape
    out = torch.empty_like(x)
    if B == 0 or L == 0:
        return out
    L_ = lib()
    chunks = int(L_.sbfe_wav_abs_chunks(L))
    partials = torch.empty((2, B, chunks), dtype=torch.float32, device=x.device)
    with _on_device(x.device):
        check(L_.sbfe_wav_add_babble(_ptr(x), B, L, int(speakers), _ptr(partials), _ptr(lengths),
                                     _ptr(snr_db), _ptr(out), _stream()))
    return out


def wav_resample(x, weights, first, in_stride, n_out):
    """Polyphase resampling of x [B, L] -> [B, n_out] (sbfe_wav_resample); weights [P, W] float32
    and first [P] int32 on x's device."""
    B, L = x.shape
    out = torch.empty((B, n_out), dtype=torch.float32, device=x.device)
    P, W = weights.shape
    with _on_device(x.device):
        check(lib().sbfe_wav_resample(_ptr(x), B, L, _ptr(weights), _ptr(first), P, int(in_stride), W,
                                      n_out, _ptr(out), _stream()))
    return out
