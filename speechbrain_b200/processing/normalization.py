"""``InputNormalization`` drop-in (reference: speechbrain/processing/features.py:940-1231).

The reference loops over sentences in Python (one host sync and >= 5 tiny kernels
per sentence).  Here the statistics of the whole batch come from one or two kernel
launches (one tiny one when the features come straight from this package's Fbank,
whose kernel leaves per-tile moments behind) and the normalisation from one; nothing
is read back to the host except in ``speaker`` mode, whose dictionary is keyed by
host integers in the reference as well.
"""
import torch

from .. import ops
from .._jit import traceable

__all__ = ["InputNormalization"]


def _dist_ready():
    return torch.distributed.is_available() and torch.distributed.is_initialized()


def sync_payload(cur, group=None):
    """All-reduce (sum) of the ``[2F+1]`` statistics payload ``{sum_b mean_b, sum_b std_b, B}``:
    the one collective of the path.  Statistics are means of per-sentence
    vectors, so summing payloads over ranks and dividing by the summed count equals the
    reference run single-process on the concatenated global batch, uneven shards included.
    NCCL on the GPUs of a job; device-agnostic so the host logic is testable over gloo."""
    if _dist_ready() and torch.distributed.get_world_size(group) > 1:
        torch.distributed.all_reduce(cur, op=torch.distributed.ReduceOp.SUM, group=group)
    return cur


def _swappable(x):
    """x owns its whole storage and nothing tracks it: x.set_(y) is then indistinguishable from
    an in-place write for whoever holds x (views of x made earlier keep the old values)."""
    return (x._base is None and not x.requires_grad and x.storage_offset() == 0
            and x.untyped_storage().nbytes() == x.numel() * x.element_size()
            and (not x.is_inference() or torch.is_inference_mode_enabled()))


class InputNormalization(torch.nn.Module):
    """Mean / variance normalisation: ``sentence``, ``speaker``, ``batch`` or
    ``global`` statistics, same arguments and state as the reference.

    Extra (not in the reference): ``sync_stats``.  When True and a
    ``torch.distributed`` group is initialised, ``batch``/``global`` statistics
    are all-reduced (2F+1 floats) so every rank normalises with the statistics
    of the concatenated global batch.  Default False = the reference's per-rank
    behaviour.
    """

    from typing import Dict

    spk_dict_mean: Dict[int, torch.Tensor]
    spk_dict_std: Dict[int, torch.Tensor]
    spk_dict_count: Dict[int, int]

    def __init__(self, mean_norm=True, std_norm=True, norm_type="global", avg_factor=None,
                 requires_grad=False, update_until_epoch=3, sync_stats=False, output_dtype=None):
        super().__init__()
        # not in the reference (opt-in, like sync_stats): torch.bfloat16 makes forward return a
        # NEW bfloat16 tensor (the float32 result rounded to nearest even: relative error <= 2^-8
        # per value; statistics stay float32) - half the bytes for the model's first layer and
        # for a device -> host copy.  The in-place modes then leave x untouched.
        if output_dtype not in (None, torch.float32, torch.bfloat16):
            raise ValueError("output_dtype must be None, torch.float32 or torch.bfloat16")
        self.output_dtype = None if output_dtype == torch.float32 else output_dtype
        register_checkpoint_hooks_if_available(import_it=False)
        self.mean_norm = mean_norm
        self.std_norm = std_norm
        self.norm_type = norm_type
        self.avg_factor = avg_factor
        self.requires_grad = requires_grad
        self.glob_mean = torch.tensor([0])
        self.glob_std = torch.tensor([0])
        self.spk_dict_mean = {}
        self.spk_dict_std = {}
        self.spk_dict_count = {}
        self.weight = 1.0
        self.count = 0
        self.eps = 1e-10
        self.update_until_epoch = update_until_epoch
        self.sync_stats = sync_stats
        self.device_inp = torch.device("cpu")
        # speaker mode: device tables behind spk_dict_mean / spk_dict_std (see _speaker_table)
        self._spk_tab, self._spk_row, self._spk_dirty = None, {}, False

    # ------------------------------------------------------------------ forward
    @traceable(lambda self: 0 if self.norm_type in ("sentence", "speaker") else None)
    def forward(self, x, lengths, spk_ids=torch.tensor([]), epoch=0):
        """x ``[B, T, F]``, lengths ``[B]`` relative; see :1000-1118."""
        if x.dim() != 3:
            raise NotImplementedError("InputNormalization takes [batch, time, features]")
        if x.dtype != torch.float32 or not x.is_contiguous():
            if self.norm_type in ("sentence", "speaker"):
                raise TypeError("in-place normalisation needs a contiguous float32 tensor")
            x = x.float().contiguous()
        ops._require_cuda(x, "x")
        B, T, F = x.shape
        # one-shot hand-over from the fused Fbank kernel (lobes/features.py): tile records of
        # this very tensor, void as soon as anything has written to it (version counter; the
        # package's own in-place ops bump it, ops.mark_written)
        ts = getattr(x, "_sbfe_tile_stats", None)
        if ts is not None:
            x._sbfe_tile_stats = None
            if (ops.version_of(x) != ts.version or ts.version is None
                    or tuple(ts.stats.shape[::2]) != (B, 3 * F + 4)):
                ts = None

        if self.norm_type == "sentence" and ts is not None:
            # statistics + normalisation in one library call, in place (:1020-1035)
            # the version check above is the statement SBFE_BESIDE_CLAMP asks for: nothing this
            # package or torch launched has written x since Fbank returned it (same stream)
            beside = ts.stream == ops.stream_id(x.device)
            if self.output_dtype in (None, torch.float32) and _swappable(x):
                # The kernel writes a fresh tensor and x takes over its storage: for the caller x
                # IS normalised when the call returns, like the reference's x[snt_id] = ... - but
                # no CTA overwrites rows that another CTA of the utterance still reads for its
                # statistics (the tile the length cuts, clamped tiles), which the truly in-place
                # launch below has to order with tickets (csrc/norm.cu) at ~6 us per cfg2 step
                y = ops.norm_sentence_tiles(x, lengths, ts, self.mean_norm, self.std_norm, self.eps,
                                            beside_clamp=beside)
                x.set_(y)
                return x
            return ops.norm_sentence_tiles(x, lengths, ts, self.mean_norm, self.std_norm, self.eps,
                                           out=x, beside_clamp=beside, out_dtype=self.output_dtype)
        if ts is not None:
            mean, std = ops.norm_stats_tiles(x, lengths, ts, self.mean_norm, self.std_norm, self.eps)
        else:
            mean, std = ops.norm_stats(x, lengths, self.mean_norm, self.std_norm, self.eps)

        if self.norm_type == "sentence":
            return ops.norm_apply(x, mean, std, per_utt=True, out=x,  # in place (:1035)
                                  out_dtype=self.output_dtype)

        if self.norm_type == "speaker":
            s_mean, s_std = self._speaker_stats(mean, std, spk_ids)
            return ops.norm_apply(x, s_mean, s_std, per_utt=True, out=x,  # in place (:1081)
                                  out_dtype=self.output_dtype)

        if self.norm_type not in ("batch", "global"):
            return x

        cur = ops.norm_reduce(mean, std)  # {sum_b mean_b, sum_b std_b, B}
        if self.sync_stats:
            sync_payload(cur)  # the one collective of the path
        if self.norm_type == "batch":
            g_mean = torch.empty((F,), dtype=torch.float32, device=x.device)
            g_std = torch.empty((F,), dtype=torch.float32, device=x.device)
            ops.norm_update(cur, g_mean, g_std, first=True, weight=1.0)
            return ops.norm_apply(x, g_mean, g_std, per_utt=False, out_dtype=self.output_dtype)

        # global (:1090-1116)
        if self.training:
            if self.count == 0:
                self.glob_mean = torch.empty((F,), dtype=torch.float32, device=x.device)
                self.glob_std = torch.empty((F,), dtype=torch.float32, device=x.device)
                ops.norm_update(cur, self.glob_mean, self.glob_std, first=True, weight=1.0)
            elif epoch < self.update_until_epoch:
                self.weight = 1 / (self.count + 1) if self.avg_factor is None else self.avg_factor
                self._stats_to(x.device, F)
                ops.norm_update(cur, self.glob_mean, self.glob_std, first=False, weight=self.weight)
            self.count = self.count + 1
        self._stats_to(x.device, F)
        return ops.norm_apply(x, self.glob_mean, self.glob_std, per_utt=False,
                              out_dtype=self.output_dtype)

    def _stats_to(self, device, F):
        """Global stats as float32 [F] on ``device`` (the never-initialised
        ``tensor([0])`` broadcasts to zeros -> inf/nan output like the reference)."""
        for name in ("glob_mean", "glob_std"):
            t = getattr(self, name)
            if not isinstance(t, torch.Tensor):
                t = torch.tensor([t])
            if t.device != device or t.dtype != torch.float32 or t.numel() != F:
                t = t.to(device=device, dtype=torch.float32)
                if t.numel() == 1 and F != 1:
                    t = t.expand(F)
                setattr(self, name, t.contiguous())

    # ------------------------------------------------------------ speaker mode
    def _speaker_table(self, device, F, extra):
        """The per-speaker statistics as rows of two device tables; ``spk_dict_mean`` /
        ``spk_dict_std`` (the reference's attributes, also what ``_save`` writes) hold views of
        the rows, so they follow the kernel's in-place updates."""
        tab = self._spk_tab
        stale = (tab is None or tab[0].device != device or tab[0].shape[1] != F
                 or self._spk_dirty or len(self._spk_row) != len(self.spk_dict_mean))
        need = len(self.spk_dict_mean) + extra
        if stale or need > tab[0].shape[0]:
            cap = max(16, 2 * need)
            tm = torch.zeros((cap, F), dtype=torch.float32, device=device)
            ts = torch.ones((cap, F), dtype=torch.float32, device=device)
            self._spk_row = {}
            for r, spk in enumerate(self.spk_dict_mean):
                tm[r] = self.spk_dict_mean[spk].to(device=device, dtype=torch.float32)
                ts[r] = self.spk_dict_std[spk].to(device=device, dtype=torch.float32)
                self._spk_row[spk] = r
            for spk, r in self._spk_row.items():
                self.spk_dict_mean[spk], self.spk_dict_std[spk] = tm[r], ts[r]
            self._spk_tab, self._spk_dirty = (tm, ts), False
        return self._spk_tab

    def _speaker_stats(self, mean, std, spk_ids):
        """Per-speaker running statistics (:1037-1081); returns the [B, F] rows to apply.  The
        dictionary logic (ids, counts, weights) runs on the host like the reference's, the
        arithmetic in one kernel that replays the batch in order (sbfe_norm_speaker_update)."""
        B, F = mean.shape
        ids = [int(v) for v in spk_ids.reshape(B, -1)[:, 0].tolist()]
        new = len({s for s in ids if s not in self.spk_dict_mean}) if self.training else 0
        tm, ts = self._speaker_table(mean.device, F, new)
        row, mode, w, k = [], [], [], []
        fresh = {}
        for spk in ids:
            r, md, wt = -1, 0, 0.0
            if self.training:
                if spk not in self._spk_row:
                    r = self._spk_row[spk] = len(self._spk_row)
                    fresh[spk] = r
                    self.spk_dict_count[spk] = 1
                    md = 2
                else:
                    r = self._spk_row[spk]
                    self.spk_dict_count[spk] = self.spk_dict_count[spk] + 1
                    self.weight = (1 / self.spk_dict_count[spk] if self.avg_factor is None
                                   else self.avg_factor)
                    md, wt = 1, self.weight
            elif spk in self._spk_row:
                r = self._spk_row[spk]
            row.append(r)
            mode.append(md)
            w.append(wt)
            k.append(1 - wt)
        dev = mean.device
        out = ops.norm_speaker_update(
            mean, std, torch.tensor(row, dtype=torch.int32).to(dev, non_blocking=True),
            torch.tensor(mode, dtype=torch.int32).to(dev, non_blocking=True),
            torch.tensor(w + k, dtype=torch.float32).to(dev, non_blocking=True), tm, ts)
        for spk, r in fresh.items():
            self.spk_dict_mean[spk], self.spk_dict_std[spk] = tm[r], ts[r]
        return out

    # --------------------------------------------------------- state / checkpoint
    def _statistics_dict(self):
        """Same 6-key dict the reference saves (:1147-1158)."""
        return {
            "count": self.count,
            "glob_mean": self.glob_mean,
            "glob_std": self.glob_std,
            "spk_dict_mean": self.spk_dict_mean,
            "spk_dict_std": self.spk_dict_std,
            "spk_dict_count": self.spk_dict_count,
        }

    def _load_statistics_dict(self, state):
        """(:1160-1192); ``glob_mean`` may be a legacy python int."""
        self.count = state["count"]
        self.glob_mean = state["glob_mean"]
        self.glob_std = state["glob_std"]
        self.spk_dict_mean = {k: v.to(self.device_inp) for k, v in state["spk_dict_mean"].items()}
        self.spk_dict_std = {k: v.to(self.device_inp) for k, v in state["spk_dict_std"].items()}
        self.spk_dict_count = state["spk_dict_count"]
        self._spk_dirty = True
        return state

    def to(self, device):
        """Moves the statistics too (:1194-1203)."""
        self = super(InputNormalization, self).to(device)
        if isinstance(self.glob_mean, torch.Tensor):
            self.glob_mean = self.glob_mean.to(device)
            self.glob_std = self.glob_std.to(device)
        for spk in self.spk_dict_mean:
            self.spk_dict_mean[spk] = self.spk_dict_mean[spk].to(device)
            self.spk_dict_std[spk] = self.spk_dict_std[spk].to(device)
        self._spk_dirty = True
        return self

    def _save(self, path):
        """Checkpoint saver hook: ``torch.save`` of the statistics dict (:1205-1215)."""
        torch.save(self._statistics_dict(), path)

    def _load(self, path, end_of_epoch=False, device=None):
        """Checkpoint loader / transfer hook (:1217-1231)."""
        del end_of_epoch
        stats = torch.load(path, map_location=device)
        self._load_statistics_dict(stats)


def register_checkpoint_hooks_if_available(import_it=True):
    """Registers ``_save`` / ``_load`` with SpeechBrain's checkpointer (utils/checkpoints.py:
    197-328) - the reference does it with decorators at class definition (:940, 1205-1231).
    ``import_it=True`` (``speechbrain_b200.install()``) imports ``speechbrain.utils.checkpoints``;
    ``False`` (every construction, and the import of this module) only looks whether the
    application has imported it already, so that the class used directly with a reference
    ``Checkpointer`` - without ``install()`` - saves its statistics and not an empty
    ``state_dict``.  A no-op when the reference package is absent."""
    import sys

    ck = sys.modules.get("speechbrain.utils.checkpoints")
    if ck is None and import_it:
        try:
            from speechbrain.utils import checkpoints as ck
        except Exception:
            return False
    if ck is None:
        return False
    cls = InputNormalization
    if ck.DEFAULT_SAVE_HOOKS.get(cls) is not cls._save:
        ck.DEFAULT_SAVE_HOOKS[cls] = cls._save
        ck.DEFAULT_LOAD_HOOKS[cls] = cls._load
        ck.DEFAULT_TRANSFER_HOOKS[cls] = cls._load
    return True


register_checkpoint_hooks_if_available(import_it=False)
