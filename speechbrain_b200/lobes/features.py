"""B200-native drop-ins for ``speechbrain.lobes.features.Fbank`` / ``MFCC``.

Same constructors, sub-module names (``compute_STFT``, ``compute_fbanks``,
``compute_dct``, ``compute_deltas``, ``context_window``), ``state_dict`` keys and
``[batch, time, features]`` output as the reference (speechbrain/lobes/features.py:
22-281), but ``forward`` is one library call (two kernel launches) instead of a chain of
15-25 ATen ops:
  plain log-mel   sbfe_fbank_forward, mode (b): wav -> log-mel + per-tile records, then the
                  top_db clamp launch; the records ride on the returned tensor for a
                  following InputNormalization (processing/normalization.py)
  otherwise       sbfe_fbank_forward, mode (a): wav -> un-clamped log-mel + per-utterance max
                  sbfe_post_forward   top_db clamp [-> DCT] [-> deltas] [-> context window]
"""
import torch

from .. import ops
from .._jit import plain_eager, traceable
from ..processing.features import DCT, STFT, ContextWindow, Deltas, Filterbank

__all__ = ["Fbank", "MFCC", "pair"]


def pair(compute_features, normalizer):
    """Kept for source compatibility with round-1 recipes: a no-op.

    It used to declare that the features go straight into ``normalizer`` so that the top_db
    clamp could be deferred to it.  ``sbfe_fbank_forward`` now clamps itself (a second, small
    launch of the same call, csrc/fbank400_v3.cu) and the normalisation runs beside that launch,
    so an unmodified recipe - ``normalize(compute_features(wavs), wav_lens)`` - already gets the
    shortest chain and the tensor ``compute_features`` returns is always final."""
    if not isinstance(compute_features, _Lobe):
        raise TypeError("pair() takes a speechbrain_b200 Fbank / MFCC")
    return compute_features


class _Lobe(torch.nn.Module):
    def _make_common(self, requires_grad, sample_rate, f_min, f_max, n_fft, n_mels, filter_shape,
                     param_change_factor, param_rand_factor, left_frames, right_frames, win_length,
                     hop_length):
        if f_max is None:
            f_max = sample_rate / 2
        self.compute_STFT = STFT(sample_rate=sample_rate, n_fft=n_fft, win_length=win_length,
                                 hop_length=hop_length)
        self.compute_fbanks = Filterbank(
            sample_rate=sample_rate, n_fft=n_fft, n_mels=n_mels, f_min=f_min, f_max=f_max,
            freeze=not requires_grad, filter_shape=filter_shape,
            param_change_factor=param_change_factor, param_rand_factor=param_rand_factor)
        self.context_window = ContextWindow(left_frames=left_frames, right_frames=right_frames)
        self._tables = None

    def _frontend_tables(self):
        fb = self.compute_fbanks
        jitter = fb.param_rand_factor != 0 and fb.training
        if self._tables is None or jitter:
            mel = fb.tables()  # re-drawn per call when the reference would jitter
            st = self.compute_STFT
            st._check_supported()
            if not st._default_mode():   # the lobes build their STFT with the defaults (:107-112)
                raise NotImplementedError("Fbank / MFCC run the default STFT mode (center=True, "
                                          "pad_mode='constant', onesided=True)")
            win = st.window.detach().float().cpu()
            if st.normalized_stft:
                win = win * (float(st.n_fft) ** -0.5)
            tab = ops.FrontendTables(st.n_fft, st.hop_length, win, None, fb.log_mel, fb.multiplier,
                                     fb.amin, fb.db_multiplier, mel_precision=mel.mel_precision)
            tab.n_mels, tab._h_mel, tab.nnz = mel.n_mels, mel._h_mel, mel.nnz
            if jitter:
                return tab
            self._tables = tab
        return self._tables

    def _features(self, wav, dct):
        if not hasattr(wav, "packed") and wav.dim() != 2:
            raise NotImplementedError(
                "speechbrain_b200 Fbank/MFCC take [batch, time] waveforms (for multi-channel "
                "input the reference sums over the channel axis, lobes/features.py:138-139)")
        fb = self.compute_fbanks
        ctx = self.context and self.context_window.context_len > 1
        if not fb.freeze:
            # requires_grad=True (next row N4): learnable centres / bands.  Power spectrum from the
            # STFT kernel, then the Filterbank module's autograd-aware forward; the waveform gets
            # no gradient, and the DCT / deltas / context stages have no backward pass here
            if dct is not None or self.deltas or ctx:
                raise NotImplementedError(
                    "requires_grad=True is built for plain Fbank features (deltas=False, "
                    "context=False); MFCC / deltas / context have no backward pass")
            self.compute_STFT._check_supported()
            spec = ops.stft_forward(self.compute_STFT.tables(), ops._wav2d(wav.detach()), power_only=True)
            return fb(spec)
        if dct is None and not self.deltas and not ctx and fb.log_mel:
            # plain log-mel: ONE library call (fused kernel + its clamp launch).  A following
            # InputNormalization gets its statistics from the per-tile records handed over on the
            # returned tensor instead of a pass over it
            feats, utt_max, ts = ops.fbank_forward(self._frontend_tables(), wav, tile_stats=True,
                                                   top_db=fb.top_db)
            if ts is not None:
                ts.version = ops.version_of(feats)
                if ts.version is not None and plain_eager():
                    feats._sbfe_tile_stats = ts
                return feats
        else:
            feats, utt_max = ops.fbank_forward(self._frontend_tables(), wav)
        return ops.post_forward(
            feats, utt_max if fb.log_mel else None, fb.top_db,
            dct=dct, deltas=1 if self.deltas else 0,
            delta_window=2 * self.compute_deltas.n + 1, context=ctx,
            left=self.context_window.left_frames, right=self.context_window.right_frames,
            inplace=True)


class Fbank(_Lobe):
    """Log-mel filterbank features; reference: lobes/features.py:22-147."""

    def __init__(self, deltas=False, context=False, requires_grad=False, sample_rate=16000,
                 f_min=0, f_max=None, n_fft=400, n_mels=40, filter_shape="triangular",
                 param_change_factor=1.0, param_rand_factor=0.0, left_frames=5, right_frames=5,
                 win_length=25, hop_length=10):
        super().__init__()
        self.deltas = deltas
        self.context = context
        self.requires_grad = requires_grad
        self._make_common(requires_grad, sample_rate, f_min, f_max, n_fft, n_mels, filter_shape,
                          param_change_factor, param_rand_factor, left_frames, right_frames,
                          win_length, hop_length)
        self.compute_deltas = Deltas(input_size=n_mels)

    @traceable()
    def forward(self, wav):
        """wav ``[B, L]`` -> ``[B, 1 + L//hop, n_mels * (3 if deltas) * (C if context)]``."""
        return self._features(wav, None)


class MFCC(_Lobe):
    """MFCC features; reference: lobes/features.py:150-281."""

    def __init__(self, deltas=True, context=True, requires_grad=False, sample_rate=16000, f_min=0,
                 f_max=None, n_fft=400, n_mels=23, n_mfcc=20, filter_shape="triangular",
                 param_change_factor=1.0, param_rand_factor=0.0, left_frames=5, right_frames=5,
                 win_length=25, hop_length=10):
        super().__init__()
        self.deltas = deltas
        self.context = context
        self.requires_grad = requires_grad
        self._make_common(requires_grad, sample_rate, f_min, f_max, n_fft, n_mels, filter_shape,
                          param_change_factor, param_rand_factor, left_frames, right_frames,
                          win_length, hop_length)
        self.compute_dct = DCT(input_size=n_mels, n_out=n_mfcc)
        self.compute_deltas = Deltas(input_size=n_mfcc)

    @traceable()
    def forward(self, wav):
        return self._features(wav, self.compute_dct.matrix_on(wav.device))
