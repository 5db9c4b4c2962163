"""Registry of golden cases shared by ``make_golden.py`` (runs the unmodified
reference), ``test_oracle_golden.py`` (numpy oracle) and the ``-m gpu`` parity
tests (CUDA product).  Inputs are regenerated from seeds; the committed
``tests/golden/frontend_golden.npz`` holds the reference outputs plus a checksum
of every regenerated input so a drifting RNG is caught rather than mis-compared.
"""
import os

import numpy as np

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
GOLDEN_NPZ = os.path.join(GOLDEN_DIR, "frontend_golden.npz")
SPEECH_NPZ = os.path.join(GOLDEN_DIR, "speech_excerpt.npz")


def _rng(seed):
    return np.random.Generator(np.random.PCG64(seed))


def synth_wav(seed, batch, n_samples, scale=0.1, rel_lengths=None):
    """Synthetic 16 kHz batch: gaussian noise, zero-padded past each utterance's
    relative length (mirrors ``batch_pad_right``, utils/data_utils.py:391)."""
    wav = (_rng(seed).standard_normal((batch, n_samples)) * scale).astype(np.float32)
    if rel_lengths is not None:
        for i, r in enumerate(rel_lengths):
            wav[i, int(round(float(r) * n_samples)):] = 0.0
    return wav


def synth_feats(seed, shape, scale=3.0, offset=-20.0):
    return (_rng(seed).standard_normal(shape) * scale + offset).astype(np.float32)


def speech_wav():
    """1.5 s of real speech (int16 PCM stored in the fixture, 16 kHz)."""
    pcm = np.load(SPEECH_NPZ)["pcm"]
    return (pcm.astype(np.float32) / 32768.0)[None, :]


# Each case: name -> dict(kind=..., make_input=callable, kwargs=..., extras)
# kinds: fbank, mfcc, fbank+norm, stft, magnitude, filterbank, dct, deltas,
#        context, norm_seq (a sequence of InputNormalization calls).
CASES = {}


def _case(name, **spec):
    CASES[name] = spec


LENS4 = [1.0, 0.8, 0.6, 0.9]

_case("cfg1_fbank40_8x1s", kind="fbank", make_input=lambda: synth_wav(101, 8, 16000),
      kwargs={})
_case("cfg2_fbank80_sentnorm_4x2s", kind="fbank+norm",
      make_input=lambda: synth_wav(102, 4, 32000, rel_lengths=LENS4),
      kwargs={"n_mels": 80}, norm={"norm_type": "sentence"}, lengths=LENS4)
_case("cfg3_mfcc660_2xhalf", kind="mfcc", make_input=lambda: synth_wav(103, 2, 8000),
      kwargs={})
_case("cfg4_fbank80_sentmean_4x1s", kind="fbank+norm",
      make_input=lambda: synth_wav(104, 4, 16000, rel_lengths=LENS4),
      kwargs={"n_mels": 80}, norm={"norm_type": "sentence", "std_norm": False},
      lengths=LENS4)
_case("cfg5_fbank80_globalnorm_4x1s", kind="fbank+norm",
      make_input=lambda: synth_wav(105, 4, 16000, rel_lengths=LENS4),
      kwargs={"n_mels": 80}, norm={"norm_type": "global"}, lengths=LENS4)
_case("fbank40_loud_scale1", kind="fbank", make_input=lambda: synth_wav(106, 3, 8000, scale=1.0),
      kwargs={})
_case("fbank40_deltas_ctx23", kind="fbank", make_input=lambda: synth_wav(107, 2, 8000),
      kwargs={"deltas": True, "context": True, "left_frames": 2, "right_frames": 3})
_case("fbank_nfft512_64mel", kind="fbank", make_input=lambda: synth_wav(108, 2, 8000),
      kwargs={"n_fft": 512, "n_mels": 64})
_case("fbank_8k_20ms", kind="fbank", make_input=lambda: synth_wav(109, 2, 6000),
      kwargs={"sample_rate": 8000, "n_fft": 256, "win_length": 20, "hop_length": 8,
              "n_mels": 24})
_case("fbank_gaussian", kind="fbank", make_input=lambda: synth_wav(110, 2, 4000),
      kwargs={"filter_shape": "gaussian", "n_mels": 40})
_case("fbank_rectangular", kind="fbank", make_input=lambda: synth_wav(111, 2, 4000),
      kwargs={"filter_shape": "rectangular", "n_mels": 40})
_case("fbank80_speech", kind="fbank", make_input=speech_wav, kwargs={"n_mels": 80})
_case("mfcc_speech_nodelta", kind="mfcc", make_input=speech_wav,
      kwargs={"deltas": False, "context": False})
_case("fbank_short_utt", kind="fbank", make_input=lambda: synth_wav(112, 3, 100), kwargs={})
_case("fbank_silence_tail", kind="fbank",
      make_input=lambda: synth_wav(113, 2, 4800, rel_lengths=[0.25, 1.0]), kwargs={})

_case("stft_2x4000", kind="stft", make_input=lambda: synth_wav(120, 2, 4000), kwargs={"sample_rate": 16000})
_case("stft_multichannel", kind="stft",
      make_input=lambda: synth_wav(121, 6, 2400).reshape(2, 3, 2400).transpose(0, 2, 1).copy(),
      kwargs={"sample_rate": 16000})
_case("stft_nfft512", kind="stft", make_input=lambda: synth_wav(122, 2, 3000),
      kwargs={"sample_rate": 16000, "n_fft": 512})
_case("magnitude_sqrt", kind="magnitude", make_input=lambda: synth_feats(123, (2, 7, 201, 2), 1.0, 0.0),
      kwargs={"power": 0.5})
_case("magnitude_log", kind="magnitude", make_input=lambda: synth_feats(124, (2, 7, 201, 2), 1.0, 0.0),
      kwargs={"power": 1, "log": True})
_case("filterbank_default", kind="filterbank",
      make_input=lambda: np.abs(synth_feats(125, (3, 21, 201), 5.0, 0.0)), kwargs={})
_case("filterbank_nolog", kind="filterbank",
      make_input=lambda: np.abs(synth_feats(126, (2, 9, 201), 5.0, 0.0)),
      kwargs={"log_mel": False, "n_mels": 23})
_case("filterbank_amp20", kind="filterbank",
      make_input=lambda: np.abs(synth_feats(127, (2, 9, 201), 5.0, 0.0)),
      kwargs={"power_spectrogram": 1, "n_mels": 30, "top_db": 40.0})
_case("filterbank_multichannel", kind="filterbank",
      make_input=lambda: np.abs(synth_feats(128, (2, 9, 201, 3), 5.0, 0.0)), kwargs={})
_case("dct_40_20", kind="dct", make_input=lambda: synth_feats(129, (2, 11, 40)),
      kwargs={"input_size": 40})
_case("dct_23_13_noortho", kind="dct", make_input=lambda: synth_feats(130, (2, 11, 23)),
      kwargs={"input_size": 23, "n_out": 13, "ortho_norm": False})
_case("deltas_20", kind="deltas", make_input=lambda: synth_feats(131, (3, 17, 20)),
      kwargs={"input_size": 20})
_case("deltas_short_T2", kind="deltas", make_input=lambda: synth_feats(132, (2, 2, 5)),
      kwargs={"input_size": 5})
_case("deltas_win9", kind="deltas", make_input=lambda: synth_feats(133, (2, 12, 6)),
      kwargs={"input_size": 6, "window_length": 9})
_case("context_l5r5", kind="context", make_input=lambda: synth_feats(134, (2, 13, 6)),
      kwargs={"left_frames": 5, "right_frames": 5})
_case("context_l1r4", kind="context", make_input=lambda: synth_feats(135, (2, 9, 4)),
      kwargs={"left_frames": 1, "right_frames": 4})
_case("context_l3r0", kind="context", make_input=lambda: synth_feats(136, (2, 9, 4)),
      kwargs={"left_frames": 3, "right_frames": 0})

# InputNormalization call sequences: each step = (seed, lengths, epoch, training[, spk_ids])
SEQ_SHAPE = (4, 37, 12)
SEQ_LENS = [[1.0, 0.5, 0.77, 0.3], [0.9, 1.0, 0.1, 0.65], [1.0, 1.0, 1.0, 1.0],
            [0.5, 0.6, 0.7, 0.8], [0.33, 0.95, 0.5, 1.0]]


def _seq(seed0, epochs, training, spk=None):
    return [dict(seed=seed0 + i, lengths=SEQ_LENS[i], epoch=epochs[i], training=training[i],
                 spk_ids=None if spk is None else spk[i]) for i in range(len(epochs))]


_case("norm_global_seq", kind="norm_seq", shape=SEQ_SHAPE, norm={"norm_type": "global"},
      steps=_seq(200, [0, 1, 2, 3, 4], [True, True, True, True, False]))
_case("norm_global_avgfactor", kind="norm_seq", shape=SEQ_SHAPE,
      norm={"norm_type": "global", "avg_factor": 0.25},
      steps=_seq(210, [0, 0, 0, 0, 0], [True, True, True, False, True]))
_case("norm_global_meanonly", kind="norm_seq", shape=SEQ_SHAPE,
      norm={"norm_type": "global", "std_norm": False},
      steps=_seq(220, [0, 0, 5, 0, 0], [True, True, True, True, True]))
_case("norm_batch", kind="norm_seq", shape=SEQ_SHAPE, norm={"norm_type": "batch"},
      steps=_seq(230, [0, 0], [True, False]))
_case("norm_batch_stdonly", kind="norm_seq", shape=SEQ_SHAPE,
      norm={"norm_type": "batch", "mean_norm": False}, steps=_seq(240, [0], [True]))
_case("norm_sentence", kind="norm_seq", shape=SEQ_SHAPE, norm={"norm_type": "sentence"},
      steps=_seq(250, [0, 0], [True, False]))
_case("norm_sentence_nostd", kind="norm_seq", shape=SEQ_SHAPE,
      norm={"norm_type": "sentence", "std_norm": False}, steps=_seq(260, [0], [True]))
_case("norm_speaker", kind="norm_seq", shape=SEQ_SHAPE, norm={"norm_type": "speaker"},
      steps=_seq(270, [0, 0, 0, 0], [True, True, True, False],
                 spk=[[3, 5, 3, 9], [5, 5, 9, 3], [3, 3, 3, 3], [3, 7, 9, 5]]))


def input_checksum(x):
    x = np.asarray(x, dtype=np.float64)
    return np.array([x.sum(), np.abs(x).sum(), x.reshape(-1)[:: max(1, x.size // 97)].sum()])


# ---- round 2 additions: a second fixture file (frontend_golden_r2.npz, tests/make_golden_r2.py)
# so that the round-1 file keeps regenerating bit for bit -------------------------------------
GOLDEN_R2_NPZ = os.path.join(GOLDEN_DIR, "frontend_golden_r2.npz")
REF_CKPT_GLOBAL = os.path.join(GOLDEN_DIR, "ref_normalizer_global.ckpt")
REF_CKPT_SPEAKER = os.path.join(GOLDEN_DIR, "ref_normalizer_speaker.ckpt")
JITTER_SEED = 77

CASES_R2 = {}


def _case2(name, **spec):
    CASES_R2[name] = spec


# Filterbank jitter (processing/features.py:525-532): training mode, torch.manual_seed(JITTER_SEED)
# right before forward - the product draws the same torch.rand(2)
_case2("fbank_jitter_train", kind="fbank", train_seed=JITTER_SEED,
       make_input=lambda: synth_wav(301, 2, 8000),
       kwargs={"n_mels": 40, "param_rand_factor": 0.15})
_case2("mfcc_jitter_train", kind="mfcc", train_seed=JITTER_SEED + 1,
       make_input=lambda: synth_wav(302, 2, 4000),
       kwargs={"param_rand_factor": 0.3, "deltas": False, "context": False})
# n_fft = 400 with hop != 160: the scalar fused kernel (csrc/fbank400.cu)
_case2("fbank_hop128", kind="fbank", make_input=lambda: synth_wav(303, 3, 8000 + 37),
       kwargs={"n_mels": 40, "hop_length": 8})
_case2("fbank_hop200_80mel", kind="fbank", make_input=lambda: synth_wav(304, 2, 9000),
       kwargs={"n_mels": 80, "hop_length": 12.5})
# the power-of-two kernels at sizes that span several tiles
_case2("fbank_nfft512_80mel_3x1s", kind="fbank",
       make_input=lambda: synth_wav(305, 3, 16000 + 91, rel_lengths=[1.0, 0.55, 0.8]),
       kwargs={"n_fft": 512, "n_mels": 80})
_case2("fbank_nfft512_hop256", kind="fbank", make_input=lambda: synth_wav(306, 2, 12000),
       kwargs={"n_fft": 512, "n_mels": 40, "win_length": 32, "hop_length": 16})
_case2("mfcc_nfft512", kind="mfcc", make_input=lambda: synth_wav(307, 2, 6000),
       kwargs={"n_fft": 512})
# dense filters (the tensor-core contraction)
_case2("fbank_gaussian80_2x1s", kind="fbank", make_input=lambda: synth_wav(308, 2, 16000),
       kwargs={"filter_shape": "gaussian", "n_mels": 80})
_case2("fbank_gaussian_nfft512", kind="fbank", make_input=lambda: synth_wav(309, 2, 8000),
       kwargs={"filter_shape": "gaussian", "n_mels": 64, "n_fft": 512})
_case2("stft_normalized", kind="stft", make_input=lambda: synth_wav(310, 2, 4000),
       kwargs={"sample_rate": 16000, "normalized_stft": True})
_case2("stft_normalized_nfft512", kind="stft", make_input=lambda: synth_wav(311, 2, 3000),
       kwargs={"sample_rate": 16000, "n_fft": 512, "normalized_stft": True})
_case2("filterbank_gaussian_spec", kind="filterbank",
       make_input=lambda: np.abs(synth_feats(312, (3, 50, 201), 5.0, 0.0)),
       kwargs={"filter_shape": "gaussian", "n_mels": 40})

# InputNormalization state written by the REFERENCE's _save (normalizer.ckpt): the steps that
# build the state, then one eval step whose output is stored
CKPT_SHAPE = (4, 37, 12)
CKPT_STEPS = _seq(400, [0, 0, 0], [True, True, True], spk=[[3, 5, 3, 9], [5, 5, 9, 3], [3, 7, 3, 3]])
CKPT_EVAL = dict(seed=450, lengths=SEQ_LENS[3], spk_ids=[3, 11, 9, 5])
