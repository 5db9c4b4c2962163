"""B200-native drop-in for ``speechbrain.lobes.augment.SpecAugment``.

Same constructor and ``forward(x)`` as the reference (speechbrain/lobes/augment.py:32-201).
The random draws are made with the reference's own ``torch.randint`` calls in the
reference's order, so a seeded run masks / warps the same positions; the data movement
(two resamplings of the time axis + both masks) is one CUDA kernel instead of two
``interpolate`` calls, two slice assignments and two ``masked_fill_`` passes.
"""
import torch

from .. import ops
from .._jit import traceable

from ..processing.speech_augmentation import DropChunk, DropFreq, SpeedPerturb

__all__ = ["SpecAugment", "TimeDomainSpecAugment"]


class TimeDomainSpecAugment(torch.nn.Module):
    """Speed perturbation -> frequency drop -> chunk drop on ``[batch, time]`` waveforms;
    reference: lobes/augment.py:204-299 (same constructor, same order, same draws)."""

    def __init__(self, perturb_prob=1.0, drop_freq_prob=1.0, drop_chunk_prob=1.0,
                 speeds=[95, 100, 105], sample_rate=16000, drop_freq_count_low=0,
                 drop_freq_count_high=3, drop_chunk_count_low=0, drop_chunk_count_high=5,
                 drop_chunk_length_low=1000, drop_chunk_length_high=2000,
                 drop_chunk_noise_factor=0):
        super().__init__()
        self.speed_perturb = SpeedPerturb(perturb_prob=perturb_prob, orig_freq=sample_rate,
                                          speeds=speeds)
        self.drop_freq = DropFreq(drop_prob=drop_freq_prob, drop_count_low=drop_freq_count_low,
                                  drop_count_high=drop_freq_count_high)
        self.drop_chunk = DropChunk(drop_prob=drop_chunk_prob, drop_count_low=drop_chunk_count_low,
                                    drop_count_high=drop_chunk_count_high,
                                    drop_length_low=drop_chunk_length_low,
                                    drop_length_high=drop_chunk_length_high,
                                    noise_factor=drop_chunk_noise_factor)

    def forward(self, waveforms, lengths):
        with torch.no_grad():
            waveforms = self.speed_perturb(waveforms)
            waveforms = self.drop_freq(waveforms)
            waveforms = self.drop_chunk(waveforms, lengths)
        return waveforms


class SpecAugment(torch.nn.Module):
    """Time warp + frequency masks + time masks on ``[batch, time, features]`` (or
    ``[batch, channel, time, features]``); reference: lobes/augment.py:32-201.

    Like the reference, the input tensor is modified in place and returned.
    """

    def __init__(self, time_warp=True, time_warp_window=5, time_warp_mode="bicubic", freq_mask=True,
                 freq_mask_width=(0, 20), n_freq_mask=2, time_mask=True, time_mask_width=(0, 100),
                 n_time_mask=2, replace_with_zero=True):
        super().__init__()
        assert (
            time_warp or freq_mask or time_mask
        ), "at least one of time_warp, time_mask, or freq_mask should be applied"
        self.apply_time_warp = time_warp
        self.time_warp_window = time_warp_window
        self.time_warp_mode = time_warp_mode
        self.freq_mask = freq_mask
        if isinstance(freq_mask_width, int):
            freq_mask_width = (0, freq_mask_width)
        self.freq_mask_width = freq_mask_width
        self.n_freq_mask = n_freq_mask
        self.time_mask = time_mask
        if isinstance(time_mask_width, int):
            time_mask_width = (0, time_mask_width)
        self.time_mask_width = time_mask_width
        self.n_time_mask = n_time_mask
        self.replace_with_zero = replace_with_zero

    # the reference's draws, in the reference's order (:129-130, :177-183)
    def _draw_warp(self, time):
        window = self.time_warp_window
        if time - window <= window:
            return None
        c = torch.randint(window, time - window, (1,))[0]
        w = torch.randint(c - window, c + window, (1,))[0] + 1
        return int(c), int(w)

    @staticmethod
    def _draw_masks(batch, D, n_mask, width_range, device):
        mask_len = torch.randint(width_range[0], width_range[1], (batch, n_mask), device=device)
        mask_pos = torch.randint(0, max(1, D - mask_len.max()), (batch, n_mask), device=device)
        return mask_pos, mask_len

    @traceable(0)
    def forward(self, x):
        ops._require_cuda(x, "x")
        if x.dim() not in (3, 4):
            raise NotImplementedError("SpecAugment takes [batch, time, fea] or [batch, ch, time, fea]")
        if x.dtype != torch.float32 or not x.is_contiguous():
            raise TypeError("in-place augmentation needs a contiguous float32 tensor")
        flat = x.view(-1, x.shape[-2], x.shape[-1])
        batch, time, fea = flat.shape
        warp = self._draw_warp(time) if self.apply_time_warp else None
        if self.replace_with_zero:
            # one pass: warp and both masks together
            fm = (self._draw_masks(batch, fea, self.n_freq_mask, self.freq_mask_width, x.device)
                  if self.freq_mask else None)
            tm = (self._draw_masks(batch, time, self.n_time_mask, self.time_mask_width, x.device)
                  if self.time_mask else None)
            out = ops.specaugment_forward(flat, warp, self.time_warp_mode, fm, tm, 0.0)
            if out is not flat:
                flat.copy_(out)
            return x
        # fill = mean of the tensor as it is when each mask is applied (:194-196): one pass per step
        if warp is not None:
            flat.copy_(ops.specaugment_forward(flat, warp, self.time_warp_mode))
        if self.freq_mask:
            fm = self._draw_masks(batch, fea, self.n_freq_mask, self.freq_mask_width, x.device)
            ops.specaugment_forward(flat, None, freq_masks=fm, fill=ops.tensor_mean(flat))
        if self.time_mask:
            tm = self._draw_masks(batch, time, self.n_time_mask, self.time_mask_width, x.device)
            ops.specaugment_forward(flat, None, time_masks=tm, fill=ops.tensor_mean(flat))
        return x
