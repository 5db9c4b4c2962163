"""Time-domain augmentation drop-ins (reference: speechbrain/processing/speech_augmentation.py).

SURVEY.md 8f row N1, second half: the modules ``TimeDomainSpecAugment`` chains right before
``compute_features`` - ``SpeedPerturb`` / ``Resample`` (:435-820), ``DropFreq`` (:909-1010),
``DropChunk`` (:1013-1173) - and ``DoClip`` (:1176-1228).  Same constructors and the same random
draws (the reference's own ``torch.rand`` / ``torch.randint`` calls on the default generator, in
the reference's order), so a seeded run picks the same chunks, frequencies, speeds and clip value;
the data movement is one kernel each (csrc/timedomain.cu) instead of per-sentence Python loops of
slice assignments and, for the resampler, five ATen calls per polyphase branch.

Waveforms are ``[batch, time]`` float32 on a CUDA device (``[batch, time, 1]`` is accepted and
returned as such); multi-channel input raises ``NotImplementedError``.  ``AddNoise`` (:30-290)
is built for its white-noise mode (``csv_file=None``; with a CSV of noise files it raises) and
``AddBabble`` (:824-906) in full; ``AddReverb`` is not built: it sits on the reference's CSV data
loader (out of scope).
"""
import math

import torch

from .. import ops

__all__ = ["AddNoise", "AddBabble", "DropChunk", "DropFreq", "DoClip", "Resample", "SpeedPerturb"]


def _as_2d(waveforms, who):
    """[B, T] view of a [B, T] / [B, T, 1] float32 CUDA tensor, and whether to unsqueeze back."""
    ops._require_cuda(waveforms, "waveforms")
    if waveforms.dim() == 3 and waveforms.shape[2] == 1:
        return waveforms.squeeze(2), True
    if waveforms.dim() != 2:
        raise NotImplementedError("%s takes [batch, time] (or [batch, time, 1]) waveforms" % who)
    return waveforms, False


class AddNoise(torch.nn.Module):
    """Mixes white noise into the batch at a random SNR; reference :30-168 (``csv_file=None``).

    Same constructor; same draws in the same order - ``torch.rand(1)`` for ``mix_prob``,
    ``torch.rand(B, 1, device=...)`` for the SNRs, ``torch.randn_like(waveforms)`` for the noise - so
    a seeded run on the same device mixes the same noise.  The amplitude of every signal, the
    scaling, the mix and the optional peak normalisation are two or three launches
    (csrc/timedomain.cu) instead of ~15 ATen ops over the batch."""

    def __init__(self, csv_file=None, csv_keys=None, sorting="random", num_workers=0, snr_low=0,
                 snr_high=0, pad_noise=False, mix_prob=1.0, start_index=None, normalize=False,
                 replacements={}, noise_sample_rate=16000, clean_sample_rate=16000):
        super().__init__()
        if csv_file is not None:
            raise NotImplementedError(
                "speechbrain_b200.AddNoise mixes white noise (csv_file=None); noise files go "
                "through the reference's CSV data loader, which is outside this package")
        self.csv_file = csv_file
        self.csv_keys = csv_keys
        self.sorting = sorting
        self.num_workers = num_workers
        self.snr_low = snr_low
        self.snr_high = snr_high
        self.pad_noise = pad_noise
        self.mix_prob = mix_prob
        self.start_index = start_index
        self.normalize = normalize
        self.replacements = replacements

    def forward(self, waveforms, lengths):
        x, was3 = _as_2d(waveforms, "AddNoise")
        x = x.float().contiguous()
        if torch.rand(1) > self.mix_prob:                       # :131-132
            out = x.clone()
            return out.unsqueeze(2) if was3 else out
        snr = torch.rand(len(x), 1, device=x.device)            # :138-139
        snr = snr * (self.snr_high - self.snr_low) + self.snr_low
        noise = torch.randn_like(x)                             # :148
        lengths = lengths.to(device=x.device, dtype=torch.float32).contiguous()
        out = ops.wav_add_noise(x, noise, lengths, snr.reshape(-1).contiguous(), self.normalize)
        return out.unsqueeze(2) if was3 else out


class AddBabble(torch.nn.Module):
    """Babble noise from the batch's own signals; reference :824-906.  Same constructor, same
    draws (``torch.rand(1)``, ``torch.rand(B, 1, device=...)``); the rolls, sums, amplitudes and
    the mix are three launches (csrc/timedomain.cu) instead of ~25 ATen ops over the batch."""

    def __init__(self, speaker_count=3, snr_low=0, snr_high=0, mix_prob=1):
        super().__init__()
        self.speaker_count = speaker_count
        self.snr_low = snr_low
        self.snr_high = snr_high
        self.mix_prob = mix_prob

    def forward(self, waveforms, lengths):
        x, was3 = _as_2d(waveforms, "AddBabble")
        x = x.float().contiguous()
        if torch.rand(1) > self.mix_prob:                       # :881-882
            out = x.clone()
            return out.unsqueeze(2) if was3 else out
        snr = torch.rand(len(x), 1, device=x.device)            # :886-887
        snr = snr * (self.snr_high - self.snr_low) + self.snr_low
        lengths = lengths.to(device=x.device, dtype=torch.float32).contiguous()
        out = ops.wav_add_babble(x, lengths, snr.reshape(-1).contiguous(), self.speaker_count)
        return out.unsqueeze(2) if was3 else out


class DropChunk(torch.nn.Module):
    """Zeroes (or fills with white noise) random chunks of each signal; reference :1013-1173."""

    def __init__(self, drop_length_low=100, drop_length_high=1000, drop_count_low=1,
                 drop_count_high=10, drop_start=0, drop_end=None, drop_prob=1, noise_factor=0.0):
        super().__init__()
        self.drop_length_low = drop_length_low
        self.drop_length_high = drop_length_high
        self.drop_count_low = drop_count_low
        self.drop_count_high = drop_count_high
        self.drop_start = drop_start
        self.drop_end = drop_end
        self.drop_prob = drop_prob
        self.noise_factor = noise_factor
        if drop_length_low > drop_length_high:
            raise ValueError("Low limit must not be more than high limit")
        if drop_count_low > drop_count_high:
            raise ValueError("Low limit must not be more than high limit")
        if drop_end is not None and drop_end >= 0:
            if drop_start > drop_end:
                raise ValueError("Low limit must not be more than high limit")
            drop_range = drop_end - drop_start
            self.drop_length_low = min(drop_length_low, drop_range)
            self.drop_length_high = min(drop_length_high, drop_range)

    def _draw_chunks(self, h_valid, noise_draw=None):
        """This call's chunks (host; the reference's draws in the reference's order, :1120-1153):
        rounds[j] = [(utterance, start, length, noise offset), ...], the j-th chunk of every
        utterance that has one.  ``h_valid``: valid samples per utterance (python ints);
        ``noise_draw(n)`` is called once per chunk where the reference draws its noise."""
        drop_times = torch.randint(low=self.drop_count_low, high=self.drop_count_high + 1,
                                   size=(len(h_valid),)).tolist()
        rounds, noise_at = [], 0
        for i in range(len(h_valid)):
            k = drop_times[i]
            if k == 0:
                continue
            length = torch.randint(low=self.drop_length_low, high=self.drop_length_high + 1,
                                   size=(k,)).tolist()
            start_min = self.drop_start
            if start_min < 0:
                start_min += h_valid[i]
            start_max = self.drop_end
            if start_max is None:
                start_max = h_valid[i]
            if start_max < 0:
                start_max += h_valid[i]
            start_max = max(0, start_max - max(length))
            start = torch.randint(low=start_min, high=start_max + 1, size=(k,)).tolist()
            for j in range(k):
                if len(rounds) <= j:
                    rounds.append([])
                rounds[j].append((i, start[j], length[j], noise_at))
                if noise_draw is not None:
                    noise_draw(length[j])
                    noise_at += length[j]
        return rounds

    def forward(self, waveforms, lengths):
        wav, squeeze = _as_2d(waveforms, "DropChunk")
        wav = wav.float().contiguous()
        B, L = wav.shape
        out = wav.clone()
        done = (lambda t: t.unsqueeze(2)) if squeeze else (lambda t: t)
        if torch.rand(1) > self.drop_prob:
            return done(out)
        # (lengths * T).long() evaluated where the caller's lengths live, read once (the reference
        # reads lengths[i] sentence by sentence: a host sync each)
        n_valid = (lengths.to(wav.device) * L).long()
        noise = []
        rounds = self._draw_chunks(
            n_valid.cpu().tolist(),
            # the reference's device draw for this chunk (:1166)
            (lambda n: noise.append(torch.rand(n, device=wav.device))) if self.noise_factor else None)
        if not rounds:
            return done(out)
        d_noise = d_max = None
        if self.noise_factor:
            # compute_amplitude(waveforms, lengths) of signal_processing.py:25-67, average mode
            amp = torch.sum(torch.abs(wav), dim=1, keepdim=True) / n_valid.unsqueeze(1)
            d_max = (2 * amp * self.noise_factor).reshape(B).contiguous()
            d_noise = torch.cat(noise)
            groups = rounds              # later chunks overwrite earlier ones: one launch per round
        else:
            groups = [[c for r in rounds for c in r]]   # zeros commute: one launch
        for g in groups:
            ops.wav_drop_chunks(out, g, d_noise, d_max)
        return done(out)


def notch_filter(notch_freq, filter_width=101, notch_width=0.05):
    """Band-reject FIR taps ``[1, filter_width, 1]`` (host): a windowed-sinc low-pass plus a
    spectrally inverted one, float32 like signal_processing.py:373-430."""
    assert 0 < notch_freq <= 1
    assert filter_width % 2 != 0
    pad = filter_width // 2
    inputs = torch.arange(filter_width) - pad
    notch_freq += notch_width

    def sinc(x):
        mid = torch.ones(1)
        return torch.cat([torch.sin(x[:pad]) / x[:pad], mid, torch.sin(x[pad + 1:]) / x[pad + 1:]])

    window = torch.blackman_window(filter_width)
    hlpf = sinc(3 * (notch_freq - notch_width) * inputs)
    hlpf *= window
    hlpf /= torch.sum(hlpf)
    hhpf = sinc(3 * (notch_freq + notch_width) * inputs)
    hhpf *= window
    hhpf /= -torch.sum(hhpf)
    hhpf[pad] += 1
    return (hlpf + hhpf).view(1, -1, 1)


class DropFreq(torch.nn.Module):
    """Notches out random frequency bands; reference :909-1010."""

    def __init__(self, drop_freq_low=1e-14, drop_freq_high=1, drop_count_low=1, drop_count_high=2,
                 drop_width=0.05, drop_prob=1):
        super().__init__()
        self.drop_freq_low = drop_freq_low
        self.drop_freq_high = drop_freq_high
        self.drop_count_low = drop_count_low
        self.drop_count_high = drop_count_high
        self.drop_width = drop_width
        self.drop_prob = drop_prob

    def _draw_frequencies(self):
        """This call's notch frequencies (host; the reference's draws, :977-983)."""
        drop_count = torch.randint(low=self.drop_count_low, high=self.drop_count_high + 1, size=(1,))
        drop_range = self.drop_freq_high - self.drop_freq_low
        return torch.rand(drop_count) * drop_range + self.drop_freq_low

    def _draw_filter(self):
        """Draws + the composite taps built on the HOST (what the tests' reference replay uses;
        forward() builds them on the device)."""
        return self._filter_from(self._draw_frequencies())

    def _filter_from(self, drop_frequency):
        """The composite taps [filter_length] with the reference's op chain (:984-1001), host."""
        filter_length = 101
        pad = filter_length // 2
        drop_filter = torch.zeros(1, 1, filter_length)
        drop_filter[0, 0, pad] = 1
        for frequency in drop_frequency:
            notch = notch_filter(frequency, filter_length, self.drop_width).transpose(2, 1)
            drop_filter = torch.nn.functional.conv1d(drop_filter, notch, padding=pad)
        return drop_filter.reshape(-1), pad

    def forward(self, waveforms):
        wav, squeeze = _as_2d(waveforms, "DropFreq")
        wav = wav.float().contiguous()
        done = (lambda t: t.unsqueeze(2)) if squeeze else (lambda t: t)
        if torch.rand(1) > self.drop_prob:
            return done(wav.clone())
        freqs = self._draw_frequencies().tolist()
        if len(freqs) > 16:        # sbfe_wav_notch_taps takes up to 16 notches per call
            taps, pad = self._filter_from(torch.tensor(freqs))
            return done(ops.wav_fir(wav, taps.to(wav.device), pad))
        taps = ops.wav_notch_taps(freqs, self.drop_width, 101, wav.device)
        return done(ops.wav_fir(wav, taps, 50))


class DoClip(torch.nn.Module):
    """Clamps the batch at a random amplitude; reference :1176-1228."""

    def __init__(self, clip_low=0.5, clip_high=1, clip_prob=1):
        super().__init__()
        self.clip_low = clip_low
        self.clip_high = clip_high
        self.clip_prob = clip_prob

    def forward(self, waveforms):
        ops._require_cuda(waveforms, "waveforms")
        if torch.rand(1) > self.clip_prob:
            return waveforms.clone()
        clipping_range = self.clip_high - self.clip_low
        clip_value = torch.rand(1)[0] * clipping_range + self.clip_low
        return ops.wav_clip(waveforms.float().contiguous(), float(clip_value))


class Resample(torch.nn.Module):
    """Band-limited resampling (Kaldi's LinearResample); reference :511-820."""

    def __init__(self, orig_freq=16000, new_freq=16000, lowpass_filter_width=6):
        super().__init__()
        self.orig_freq = orig_freq
        self.new_freq = new_freq
        self.lowpass_filter_width = lowpass_filter_width
        base_freq = math.gcd(self.orig_freq, self.new_freq)       # _compute_strides :567-580
        self.output_samples = self.new_freq // base_freq
        self.conv_stride = self.orig_freq // base_freq
        self.conv_transpose_stride = self.output_samples
        assert self.orig_freq % self.conv_stride == 0
        assert self.new_freq % self.conv_transpose_stride == 0
        self._dev = {}

    def _output_samples(self, input_num_samp):
        """LinearResample::GetNumOutputSamples (:696-734)."""
        samp_in, samp_out = int(self.orig_freq), int(self.new_freq)
        tick_freq = abs(samp_in * samp_out) // math.gcd(samp_in, samp_out)
        interval_length = input_num_samp * (tick_freq // samp_in)
        if interval_length <= 0:
            return 0
        ticks_per_output_period = tick_freq // samp_out
        last_output_samp = interval_length // ticks_per_output_period
        if last_output_samp * ticks_per_output_period == interval_length:
            last_output_samp -= 1
        return last_output_samp + 1

    def _indices_and_weights(self):
        """LinearResample::SetIndexesAndWeights in float32 on the host (:736-820): a Hann-windowed
        sinc per output phase."""
        min_freq = min(self.orig_freq, self.new_freq)
        lowpass_cutoff = 0.99 * 0.5 * min_freq
        assert lowpass_cutoff * 2 <= min_freq
        window_width = self.lowpass_filter_width / (2.0 * lowpass_cutoff)
        output_t = torch.arange(start=0.0, end=self.output_samples)
        output_t /= self.new_freq
        min_t = output_t - window_width
        max_t = output_t + window_width
        min_input_index = torch.ceil(min_t * self.orig_freq)
        max_input_index = torch.floor(max_t * self.orig_freq)
        num_indices = max_input_index - min_input_index + 1
        j = torch.arange(num_indices.max())
        input_index = min_input_index.unsqueeze(1) + j.unsqueeze(0)
        delta_t = (input_index / self.orig_freq) - output_t.unsqueeze(1)
        weights = torch.zeros_like(delta_t)
        inside = delta_t.abs().lt(window_width)
        weights[inside] = 0.5 * (1 + torch.cos(
            2 * math.pi * lowpass_cutoff / self.lowpass_filter_width * delta_t[inside]))
        zero = delta_t.eq(0.0)
        weights[~zero] *= torch.sin(2 * math.pi * lowpass_cutoff * delta_t[~zero]) / (
            math.pi * delta_t[~zero])
        weights[zero] *= 2 * lowpass_cutoff
        weights /= self.orig_freq
        self.first_indices = min_input_index
        self.weights = weights

    def forward(self, waveforms):
        if not hasattr(self, "first_indices"):
            self._indices_and_weights()
        if self.orig_freq == self.new_freq:
            return waveforms
        wav, squeeze = _as_2d(waveforms, "Resample")
        wav = wav.float().contiguous()
        tab = self._dev.get(wav.device)
        if tab is None:
            tab = (self.weights.float().contiguous().to(wav.device),
                   self.first_indices.to(torch.int32).to(wav.device))
            self._dev[wav.device] = tab
        out = ops.wav_resample(wav, tab[0], tab[1], self.conv_stride,
                               self._output_samples(wav.shape[1]))
        return out.unsqueeze(2) if squeeze else out


class SpeedPerturb(torch.nn.Module):
    """Resamples the batch to one of ``speeds`` percent of its rate; reference :435-508."""

    def __init__(self, orig_freq, speeds=[90, 100, 110], perturb_prob=1.0):
        super().__init__()
        self.orig_freq = orig_freq
        self.speeds = speeds
        self.perturb_prob = perturb_prob
        self.samp_index = 0
        self.resamplers = [Resample(orig_freq=self.orig_freq, new_freq=self.orig_freq * s // 100)
                           for s in self.speeds]

    def forward(self, waveform):
        if torch.rand(1) > self.perturb_prob:
            return waveform.clone()
        self.samp_index = torch.randint(len(self.speeds), (1,))[0]
        return self.resamplers[self.samp_index](waveform)
