"""Ragged batch ingest: the step before the feature frontend.

The reference pads on the CPU (``PaddedBatch`` / ``batch_pad_right``,
speechbrain/dataio/batch.py:20-190, speechbrain/utils/data_utils.py:347-453) and then copies
the padding over PCIe with ``batch.to(device)``.  ``RaggedBatch`` keeps only the valid samples
in ONE pinned host buffer, uploads that, and hands the fused Fbank kernel per-utterance
offsets; the kernel's copy engine zero-fills the padding on the fly, so the padded ``[B, L]``
tensor is never materialised and the features are bit-identical to the padded path.

16-bit PCM (what the audio files hold; the reference's ``read_audio`` converts to float32 on
the CPU, speechbrain/dataio/dataio.py, and then ships 4 bytes per sample) can stay int16 until
it is on the device: ``RaggedBatch.from_list`` of int16 tensors keeps them packed as int16 and
the fused kernel converts while it stages the samples (``scale`` = 1/32768, exact), halving
the PCIe bytes again; results are bit-identical to the float32 route.
"""
import torch

__all__ = ["RaggedBatch"]

_ALIGN = 4  # samples: 16-byte aligned utterance starts


class RaggedBatch:
    """Packed variable-length waveforms.

    Attributes: ``packed`` float32 (or int16 PCM, value = ``scale`` * sample) ``[n]``,
    ``offsets`` int64 ``[B]``, ``nsamples`` int32
    ``[B]`` (all on one device), ``max_len`` = L of the equivalent padded batch, ``lengths``
    float32 ``[B]`` relative lengths (what ``InputNormalization`` / the recipes call
    ``wav_lens``).
    """

    def __init__(self, packed, offsets, nsamples, max_len, lengths, scale=1.0 / 32768):
        self.packed, self.offsets, self.nsamples = packed, offsets, nsamples
        self.max_len, self.lengths = int(max_len), lengths
        self.scale = float(scale)  # only meaningful for int16 samples

    @property
    def device(self):
        return self.packed.device

    @property
    def batch_size(self):
        return int(self.offsets.numel())

    @property
    def nbytes(self):
        """Bytes that cross PCIe when this batch is uploaded."""
        return (self.packed.numel() * self.packed.element_size() + self.offsets.numel() * 8 + self.nsamples.numel() * 4
                + self.lengths.numel() * 4)

    @staticmethod
    def layout(sizes):
        """(offsets, total) of the packed buffer for utterance sizes ``sizes``."""
        offsets, fill = [], 0
        for n in sizes:
            offsets.append(fill)
            fill += (int(n) + _ALIGN - 1) // _ALIGN * _ALIGN
        return offsets, max(fill, _ALIGN)

    @classmethod
    def from_list(cls, wavs, pin_memory=True):
        """Host-side collate: list of 1-D float (or all-int16 PCM) tensors -> packed batch in
        (pinned) host memory."""
        sizes = [int(w.numel()) for w in wavs]
        offsets, total = cls.layout(sizes)
        pcm = len(wavs) > 0 and all(w.dtype == torch.int16 for w in wavs)
        packed = torch.zeros(total, dtype=torch.int16 if pcm else torch.float32, pin_memory=pin_memory)
        for w, o, n in zip(wavs, offsets, sizes):
            packed[o:o + n] = w.reshape(-1)
        max_len = max(sizes) if sizes else 0
        mk = (lambda *a, **k: torch.tensor(*a, **k).pin_memory()) if pin_memory else torch.tensor
        return cls(packed, mk(offsets, dtype=torch.int64), mk(sizes, dtype=torch.int32), max_len,
                   mk([n / max(max_len, 1) for n in sizes], dtype=torch.float32))

    @classmethod
    def from_padded(cls, wav, lengths, pin_memory=True):
        """``[B, L]`` padded host batch + relative lengths -> packed (what a collate would give)."""
        L = wav.shape[1]
        sizes = [int(round(float(r) * L)) for r in lengths]
        rb = cls.from_list([wav[i, :n] for i, n in enumerate(sizes)], pin_memory=pin_memory)
        rb.max_len = L
        rb.lengths = lengths.to(torch.float32).clone()
        if pin_memory:
            rb.lengths = rb.lengths.pin_memory()
        return rb

    def to(self, device, non_blocking=True):
        """Upload (one large + three tiny copies)."""
        mv = lambda t: t.to(device, non_blocking=non_blocking)  # noqa: E731
        return RaggedBatch(mv(self.packed), mv(self.offsets), mv(self.nsamples), self.max_len,
                           mv(self.lengths), self.scale)

    def to_padded(self):
        """Materialise the zero-padded ``[B, L]`` tensor (configurations without a ragged kernel,
        tests): one gather on the batch's device, nothing is read back to the host."""
        idx = torch.arange(self.max_len, device=self.device).unsqueeze(0)
        live = idx < self.nsamples.to(torch.int64).unsqueeze(1)
        src = (self.offsets.unsqueeze(1) + idx).clamp_(max=max(self.packed.numel() - 1, 0))
        out = torch.where(live, self.packed[src].to(torch.float32), 0.0)   # int16 -> float32 is exact
        return out * self.scale if self.packed.dtype == torch.int16 else out
