"""B200-native drop-ins for ``speechbrain.processing.features``.

Same class names, constructor signatures, defaults, attributes and
``[batch, time, features]`` layouts as the reference
(speechbrain/processing/features.py); the arithmetic runs in hand-written
sm_100a kernels (speechbrain_b200/csrc) behind the C ABI of include/sbfe.h.
CUDA tensors only - CPU input raises (the reference stays the CPU path).
"""
import logging
import math

import torch

from .. import ops
from .._jit import traceable

logger = logging.getLogger(__name__)

__all__ = ["STFT", "ISTFT", "spectral_magnitude", "Filterbank", "DCT", "Deltas", "ContextWindow",
           "InputNormalization"]


class STFT(torch.nn.Module):
    """Short-term Fourier transform; reference: processing/features.py:50-188.

    ``[B, L]`` -> ``[B, T, n_fft//2+1, 2]``; ``[B, L, C]`` -> ``[B, T, F, 2, C]``.
    ``center`` True / False, ``pad_mode`` "constant" / "reflect" and ``onesided`` True / False are
    all native (torch.stft's other pad modes, "replicate" / "circular", raise).
    """

    def __init__(self, sample_rate, win_length=25, hop_length=10, n_fft=400,
                 window_fn=torch.hamming_window, normalized_stft=False, center=True,
                 pad_mode="constant", onesided=True):
        super().__init__()
        self.sample_rate = sample_rate
        self.n_fft = n_fft
        self.normalized_stft = normalized_stft
        self.center = center
        self.pad_mode = pad_mode
        self.onesided = onesided
        # ms -> samples (processing/features.py:124-129)
        self.win_length = int(round((self.sample_rate / 1000.0) * win_length))
        self.hop_length = int(round((self.sample_rate / 1000.0) * hop_length))
        self.window = window_fn(self.win_length)
        self._tables = None

    def _check_supported(self):
        if self.pad_mode not in ("constant", "reflect"):
            raise NotImplementedError(
                "speechbrain_b200.STFT implements pad_mode 'constant' and 'reflect' (got %r)"
                % (self.pad_mode,))

    def _default_mode(self):
        """The combination every recipe uses (and the fused Fbank / MFCC kernels implement)."""
        return bool(self.center) and self.pad_mode == "constant" and bool(self.onesided)

    def tables(self):
        if self._tables is None:
            win = self.window.detach().float().cpu()
            if self.normalized_stft:  # torch.stft(normalized=True) scales by n_fft**-0.5
                win = win * (float(self.n_fft) ** -0.5)
            self._tables = ops.FrontendTables(self.n_fft, self.hop_length, win)
        return self._tables

    @traceable()
    def forward(self, x):
        self._check_supported()
        if x.dim() == 3:  # [B, L, C] -> [B*C, L]   (processing/features.py:143-146)
            b, l, c = x.shape
            flat = x.transpose(1, 2).reshape(b * c, l)
            out = ops.stft_forward(self.tables(), flat, **self._mode())  # [B*C, T, F, 2]
            out = out.reshape(b, c, out.shape[1], out.shape[2], 2)
            return out.permute(0, 2, 3, 4, 1)  # [B, T, F, 2, C]    (:175-183)
        return ops.stft_forward(self.tables(), x, **self._mode())

    def _mode(self):
        return {"center": bool(self.center), "reflect": self.pad_mode == "reflect",
                "onesided": bool(self.onesided)}


class ISTFT(torch.nn.Module):
    """Inverse short-term Fourier transform; reference: processing/features.py:191-324
    (``torch.istft``).  ``[B, T, F, 2]`` -> ``[B, L]``; ``[B, T, F, 2, C]`` -> ``[B, L, C]``.
    ``L = sig_length`` or ``(T - 1) * hop``.  Like the reference, ``normalized_stft`` and
    ``epsilon`` are accepted and never reach the transform (:309-318).  Supported natively:
    ``center=True/False`` with ``onesided=True``."""

    def __init__(self, sample_rate, n_fft=None, win_length=25, hop_length=10,
                 window_fn=torch.hamming_window, normalized_stft=False, center=True,
                 onesided=True, epsilon=1e-12):
        super().__init__()
        self.sample_rate = sample_rate
        self.n_fft = n_fft
        self.normalized_stft = normalized_stft
        self.center = center
        self.onesided = onesided
        self.epsilon = epsilon
        self.win_length = int(round((self.sample_rate / 1000.0) * win_length))
        self.hop_length = int(round((self.sample_rate / 1000.0) * hop_length))
        self.window = window_fn(self.win_length)
        self._cache = {}

    def _window_on(self, n_fft, device):
        """The window centred inside n_fft (torch.istft pads it the same way), per device."""
        key = (n_fft, device.type, device.index)
        hit = self._cache.get(key)
        if hit is None:
            win = self.window.detach().to("cpu", torch.float32)
            if win.numel() > n_fft:
                raise ValueError("win_length (%d samples) > n_fft (%d)" % (win.numel(), n_fft))
            left = (n_fft - win.numel()) // 2
            win = torch.nn.functional.pad(win, (left, n_fft - win.numel() - left)).contiguous()
            hit = self._cache[key] = (win, win.to(device))
        return hit

    def _check_envelope(self, h_win, n_fft, frames, start, out_len):
        """torch.istft refuses a window whose overlap-add vanishes somewhere in the output."""
        key = ("env", n_fft, frames, start, out_len)
        if key not in self._cache:
            full = n_fft + self.hop_length * (frames - 1)
            # overlap-add of the squared window = transposed convolution of an impulse train
            env = torch.nn.functional.conv_transpose1d(
                torch.ones(1, 1, frames), (h_win * h_win).view(1, 1, n_fft), stride=self.hop_length).view(-1)
            seg = env[start:min(start + out_len, full)]
            self._cache[key] = bool(seg.numel() == 0 or float(seg.abs().min()) >= 1e-11)
        if not self._cache[key]:
            raise RuntimeError("istft: window overlap add min is below 1e-11 (same check as torch.istft)")

    def forward(self, x, sig_length=None):
        if not self.onesided:
            raise NotImplementedError("speechbrain_b200.ISTFT implements onesided=True")
        or_shape = x.shape
        n_fft = (x.shape[2] - 1) * 2 if self.n_fft is None else self.n_fft
        if len(or_shape) == 5:  # [B, T, F, 2, C] -> [B*C, T, F, 2]   (:300-305)
            x = x.permute(0, 4, 1, 2, 3).reshape(-1, or_shape[1], or_shape[2], 2)
        elif len(or_shape) != 4:
            raise ValueError("expected [batch, time, freq, 2(, channels)], got %s" % (tuple(or_shape),))
        ops._require_cuda(x, "x")
        frames = x.shape[1]
        start = n_fft // 2 if self.center else 0
        full = n_fft + self.hop_length * (frames - 1)
        out_len = int(sig_length) if sig_length is not None else (full - 2 * start if self.center else full)
        h_win, d_win = self._window_on(n_fft, x.device)
        self._check_envelope(h_win, n_fft, frames, start, out_len)
        out = ops.istft_forward(x, n_fft, self.hop_length, d_win, self.center, out_len)
        if len(or_shape) == 5:  # back to [B, L, C]   (:320-322)
            out = out.reshape(or_shape[0], or_shape[4], -1).transpose(1, 2)
        return out


def spectral_magnitude(stft, power=1, log=False, eps=1e-14):
    """``(re^2 + im^2) ** power``; reference: processing/features.py:327-356."""
    return ops.spectral_magnitude(stft, power=power, log=log, eps=eps)


class Filterbank(torch.nn.Module):
    """Mel filterbank + amplitude_to_DB; reference: processing/features.py:359-712."""

    def __init__(self, n_mels=40, log_mel=True, filter_shape="triangular", f_min=0, f_max=8000,
                 n_fft=400, sample_rate=16000, power_spectrogram=2, amin=1e-10, ref_value=1.0,
                 top_db=80.0, param_change_factor=1.0, param_rand_factor=0.0, freeze=True,
                 mel_precision="fp32"):
        super().__init__()
        # not in the reference: dense (gaussian) filterbanks run as a tensor-core GEMM;
        # "fp32" = 3xTF32 split (inside the fp32 tolerance), "tf32" = one TF32 pass (looser
        # stated bound, include/sbfe.h: mel_precision).  May be changed before the first call.
        if mel_precision not in ("fp32", "tf32"):
            raise ValueError("mel_precision must be 'fp32' or 'tf32'")
        self.mel_precision = mel_precision
        self.n_mels = n_mels
        self.log_mel = log_mel
        self.filter_shape = filter_shape
        self.f_min = f_min
        self.f_max = f_max
        self.n_fft = n_fft
        self.sample_rate = sample_rate
        self.power_spectrogram = power_spectrogram
        self.amin = amin
        self.ref_value = ref_value
        self.top_db = top_db
        self.freeze = freeze
        self.n_stft = self.n_fft // 2 + 1
        self.db_multiplier = math.log10(max(self.amin, self.ref_value))
        self.device_inp = torch.device("cpu")
        self.param_change_factor = param_change_factor
        self.param_rand_factor = param_rand_factor
        self.multiplier = 10 if self.power_spectrogram == 2 else 20
        if self.f_min >= self.f_max:  # the reference only logs this (:457-462)
            logger.error("Require f_min: %f < f_max: %f", self.f_min, self.f_max)
        # mel-spaced centres and bands in float32, as the reference builds them (:464-488)
        mel = torch.linspace(self._to_mel(self.f_min), self._to_mel(self.f_max), self.n_mels + 2)
        hz = self._to_hz(mel)
        band = hz[1:] - hz[:-1]
        self.band = band[:-1]
        self.f_central = hz[1:-1]
        if not self.freeze:  # learnable centres / bands, stored scaled like the reference (:476-482)
            self.f_central = torch.nn.Parameter(
                self.f_central / (self.sample_rate * self.param_change_factor))
            self.band = torch.nn.Parameter(self.band / (self.sample_rate * self.param_change_factor))
        all_freqs = torch.linspace(0, self.sample_rate // 2, self.n_stft)
        self.all_freqs_mat = all_freqs.repeat(self.f_central.shape[0], 1)
        self._tables = None
        self._identity = {}

    @staticmethod
    def _to_mel(hz):
        return 2595 * math.log10(1 + hz / 700)

    @staticmethod
    def _to_hz(mel):
        return 700 * (10 ** (mel / 2595) - 1)

    def _create_fbank_matrix(self, f_central_mat, band_mat):
        """``[n_stft, n_mels]`` weights (host, float32); shapes per :586-658."""
        freqs = self.all_freqs_mat
        if self.filter_shape == "triangular":
            slope = (freqs - f_central_mat) / band_mat
            tri = torch.minimum(slope + 1.0, 1.0 - slope)
            return torch.clamp(tri, min=0.0).t()
        if self.filter_shape == "rectangular":
            inside = freqs.ge(f_central_mat - band_mat) & freqs.le(f_central_mat + band_mat)
            return inside.float().t()
        return torch.exp(-0.5 * ((freqs - f_central_mat) / (band_mat / 2.0)) ** 2).t()

    def _host_params(self):
        """f_central / band in Hz on the host (a learnable filterbank stores them scaled)."""
        fc, bw = self.f_central.detach().float().cpu(), self.band.detach().float().cpu()
        if not self.freeze:
            k = self.sample_rate * self.param_change_factor * self.param_change_factor
            fc, bw = fc * k, bw * k
        return fc, bw

    def _build_tables(self, jitter=None):
        fc, bw = self._host_params()
        fc = fc.unsqueeze(1).expand(-1, self.n_stft)
        bw = bw.unsqueeze(1).expand(-1, self.n_stft)
        if jitter is not None:
            fc, bw = fc * jitter[0], bw * jitter[1]
        mat = self._create_fbank_matrix(fc, bw)
        return ops.FrontendTables(self.n_fft, 1, torch.ones(self.n_fft), mat, self.log_mel,
                                  self.multiplier, self.amin, self.db_multiplier,
                                  mel_precision=1 if self.mel_precision == "tf32" else 0)

    def tables(self):
        """Tables for this call: cached, or re-drawn when the reference would
        jitter the filters (param_rand_factor != 0 in training mode, :525-532)."""
        if self.param_rand_factor != 0 and self.training:
            jitter = 1.0 + torch.rand(2) * 2 * self.param_rand_factor - self.param_rand_factor
            return self._build_tables(jitter)
        if self._tables is None:
            self._tables = self._build_tables()
        return self._tables

    def fbank_matrix(self):
        """Dense ``[n_stft, n_mels]`` matrix (host), for inspection/tests."""
        fc, bw = self._host_params()
        return self._create_fbank_matrix(fc.unsqueeze(1).expand(-1, self.n_stft),
                                         bw.unsqueeze(1).expand(-1, self.n_stft))

    def _learnable_forward(self, spectrogram):
        """freeze=False (:512-522, 534-551): the matrix is rebuilt from the parameters with torch
        ops on their device (so autograd reaches f_central / band), the projection + dB + top_db
        clamp run through the kernels (dense matrix: the tensor-core projection), and
        ``_FilterbankGrad`` supplies the backward pass."""
        dev = spectrogram.device
        k = self.sample_rate * self.param_change_factor * self.param_change_factor
        fc = (self.f_central.to(dev) * k).unsqueeze(1).expand(-1, self.n_stft)
        bw = (self.band.to(dev) * k).unsqueeze(1).expand(-1, self.n_stft)
        freqs = self.all_freqs_mat.to(dev)
        if self.filter_shape == "triangular":
            # the reference's op sequence (:603-612): torch.max / torch.min split the gradient at
            # ties (a bin exactly on a filter's edge: f_min, f_max), clamp / minimum would not
            slope = (freqs - fc) / bw
            zero = torch.zeros(1, device=dev)
            mat = torch.max(zero, torch.min(slope + 1.0, -slope + 1.0)).t()
        elif self.filter_shape == "rectangular":
            mat = (freqs.ge(fc - bw) & freqs.le(fc + bw)).float().t()
        else:
            mat = torch.exp(-0.5 * ((freqs - fc) / (bw / 2.0)) ** 2).t()
        return _FilterbankGrad.apply(spectrogram, mat, self)

    @traceable()
    def forward(self, spectrogram):
        sp_shape = spectrogram.shape
        if len(sp_shape) == 4:  # [B, T, F, C] -> [B*C, T, F]     (:541-545)
            spectrogram = spectrogram.permute(0, 3, 1, 2).reshape(
                sp_shape[0] * sp_shape[3], sp_shape[1], sp_shape[2])
        if not self.freeze:
            fbanks = self._learnable_forward(spectrogram)
        else:
            fbanks, utt_max = ops.filterbank_forward(self.tables(), spectrogram)
            if self.log_mel:
                fbanks = ops.post_forward(fbanks, utt_max, self.top_db, inplace=True)
        if len(sp_shape) == 4:  # (:553-558)
            fbanks = fbanks.reshape(sp_shape[0], sp_shape[3], fbanks.shape[1], fbanks.shape[2])
            fbanks = fbanks.permute(0, 2, 3, 1)
        return fbanks

    def _amplitude_to_DB(self, x):
        """dB conversion with the per-leading-element top_db clamp (:691-712)."""
        shape = x.shape
        f = shape[-1]
        x3 = x.reshape(shape[0], -1, f)
        tab = self._identity.get(f)
        if tab is None:
            tab = ops.FrontendTables(2 * (f - 1), 1, torch.ones(max(2 * (f - 1), 0)), torch.eye(f),
                                     True, self.multiplier, self.amin, self.db_multiplier)
            self._identity[f] = tab
        x_db, utt_max = ops.filterbank_forward(tab, x3)
        return ops.post_forward(x_db, utt_max, self.top_db, inplace=True).reshape(shape)


class _FilterbankGrad(torch.autograd.Function):
    """Learnable filterbank (next row N4): y = top_db_clamp(dB(spec @ W)).

    Forward: the library's kernels on a dense device matrix (``ops.filterbank_forward_dense``:
    tensor-core projection, then the clamp).  Backward (what autograd derives for the reference's
    op chain, :548, 701-712): g_mel = g_y * multiplier / ln 10 / mel where mel > amin and y was
    not clamped; the clamp's floor = utterance maximum - top_db also carries gradient - the sum of
    g_y over an utterance's clamped elements goes to its maximum element(s), as ``amax`` /
    ``torch.max`` do; g_spec = g_mel @ W^T and g_W = spec^T @ g_mel are plain library GEMMs."""

    @staticmethod
    def forward(ctx, spec, mat, mod):
        spec = spec.float().contiguous()
        y, utt_max = ops.filterbank_forward_dense(spec, mat.detach(), mod.log_mel, mod.multiplier,
                                                  mod.amin, mod.db_multiplier)
        raw = y
        if mod.log_mel:
            y = ops.post_forward(raw, utt_max, mod.top_db)
        ctx.save_for_backward(spec, mat.detach(), raw, y)
        ctx.cfg = (mod.log_mel, float(mod.multiplier), float(mod.amin), float(mod.db_multiplier))
        return y

    @staticmethod
    def backward(ctx, g):
        spec, mat, raw, y = ctx.saved_tensors
        log_mel, mult, amin, db_mul = ctx.cfg
        g = g.contiguous()
        if log_mel:
            B = raw.shape[0]
            clamped = raw < y                                   # raised to the floor
            top = raw.reshape(B, -1).amax(dim=1).reshape(B, 1, 1)
            is_max = raw == top
            g_floor = (g * clamped).reshape(B, -1).sum(dim=1).reshape(B, 1, 1)
            g_raw = torch.where(clamped, torch.zeros_like(g), g)
            g_raw = g_raw + is_max * (g_floor / is_max.reshape(B, -1).sum(dim=1).reshape(B, 1, 1))
            # raw = mult * log10(max(mel, amin)) - mult * db_mul  ->  mel from raw where mel > amin
            mel = torch.pow(10.0, (raw + mult * db_mul) / mult)
            above = raw > (mult * math.log10(amin) - mult * db_mul)
            g_mel = torch.where(above, g_raw * (mult / math.log(10.0)) / mel, torch.zeros_like(g))
        else:
            g_mel = g
        g_spec = torch.matmul(g_mel, mat.t()) if ctx.needs_input_grad[0] else None
        g_mat = None
        if ctx.needs_input_grad[1]:
            g_mat = torch.matmul(spec.reshape(-1, spec.shape[-1]).t(), g_mel.reshape(-1, g_mel.shape[-1]))
        return g_spec, g_mat, None


class DCT(torch.nn.Module):
    """DCT-II of the last axis; reference: processing/features.py:715-786."""

    def __init__(self, input_size, n_out=20, ortho_norm=True):
        super().__init__()
        if n_out > input_size:
            raise ValueError(
                "Cannot select more DCT coefficients than inputs "
                "(n_out=%i, n_in=%i)" % (n_out, input_size))
        n = torch.arange(float(input_size))
        k = torch.arange(float(n_out)).unsqueeze(1)
        dct = torch.cos(math.pi / float(input_size) * (n + 0.5) * k)
        if ortho_norm:
            dct[0] *= 1.0 / math.sqrt(2.0)
            dct *= math.sqrt(2.0 / float(input_size))
        else:
            dct *= 2.0
        self.dct_mat = dct.t().contiguous()
        self._dev = {}

    def matrix_on(self, device):
        key = (device.type, device.index)
        if key not in self._dev:
            self._dev[key] = self.dct_mat.float().contiguous().to(device)
        return self._dev[key]

    @traceable()
    def forward(self, x):
        input_shape = x.shape
        if len(input_shape) == 4:  # same flat reshape as the reference (:774-775)
            x = x.reshape(x.shape[0] * x.shape[3], x.shape[1], x.shape[2])
        out = ops.post_forward(x, dct=self.matrix_on(x.device))
        if len(input_shape) == 4:
            out = out.reshape(input_shape[0], out.shape[1], out.shape[2], input_shape[3])
        return out


class Deltas(torch.nn.Module):
    """Regression deltas along time, replicate padding; reference:
    processing/features.py:789-852.  Keeps the ``kernel`` buffer so that
    ``state_dict()`` matches the reference (``compute_deltas.kernel``)."""

    def __init__(self, input_size, window_length=5):
        super().__init__()
        self.n = (window_length - 1) // 2
        self.denom = self.n * (self.n + 1) * (2 * self.n + 1) / 3
        self.window_length = window_length
        self.register_buffer(
            "kernel",
            torch.arange(-self.n, self.n + 1, dtype=torch.float32).repeat(input_size, 1, 1),
        )

    @traceable()
    def forward(self, x):
        if x.dim() != 3:
            raise NotImplementedError("speechbrain_b200.Deltas takes [batch, time, features]")
        return ops.post_forward(x, deltas=2, delta_window=2 * self.n + 1)


class ContextWindow(torch.nn.Module):
    """Gathers left/right context frames; reference: processing/features.py:855-937.
    ``out[b, t, f*C + c] = x[b, t + c - left, f]``, zero outside the utterance."""

    def __init__(self, left_frames=0, right_frames=0):
        super().__init__()
        self.left_frames = left_frames
        self.right_frames = right_frames
        self.context_len = self.left_frames + self.right_frames + 1
        self.kernel_len = 2 * max(self.left_frames, self.right_frames) + 1
        self.kernel = torch.eye(self.context_len, self.kernel_len)
        if self.right_frames > self.left_frames:
            self.kernel = torch.roll(self.kernel, self.right_frames - self.left_frames, 1)
        self.first_call = True

    @traceable()
    def forward(self, x):
        if x.dim() != 3:
            raise NotImplementedError("speechbrain_b200.ContextWindow takes [batch, time, features]")
        self.first_call = False
        if self.context_len == 1:
            return ops.post_forward(x)  # identity copy, like the reference's conv
        return ops.post_forward(x, context=True, left=self.left_frames, right=self.right_frames)


from .normalization import InputNormalization  # noqa: E402,F401
