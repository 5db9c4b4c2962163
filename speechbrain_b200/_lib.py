"""ctypes binding of libsbfe.so (C ABI: include/sbfe.h).

There is no CPU fallback: if the library cannot be loaded (and cannot be built
because nvcc is absent) every op raises ``SbfeUnavailable``.
"""
import ctypes as C
import os
import threading

from . import _build

__all__ = ["lib", "check", "FbankDesc", "SbfeUnavailable", "loaded_path"]


class SbfeUnavailable(RuntimeError):
    pass


class FbankDesc(C.Structure):
    """Mirror of ``sbfe_fbank_desc`` (include/sbfe.h)."""

    _fields_ = [
        ("n_fft", C.c_int32),
        ("hop", C.c_int32),
        ("n_mels", C.c_int32),
        ("log_mel", C.c_int32),
        ("multiplier", C.c_float),
        ("amin", C.c_float),
        ("db_offset", C.c_float),
        ("floor_db", C.c_float),
        ("h_window", C.c_void_p),
        ("d_window", C.c_void_p),
        ("d_mel_start", C.c_void_p),
        ("d_mel_count", C.c_void_p),
        ("d_mel_offset", C.c_void_p),
        ("d_mel_weight", C.c_void_p),
        ("mel_nnz", C.c_int32),
        ("mel_pad4", C.c_int32),
        ("mel_precision", C.c_int32),
    ]


_P, _I64, _I32, _F = C.c_void_p, C.c_int64, C.c_int32, C.c_float
_DESC = C.POINTER(FbankDesc)

# name -> (restype, argtypes); must list every symbol include/sbfe.h declares
SIGNATURES = {
    "sbfe_abi_version": (C.c_int, []),
    "sbfe_error_string": (C.c_char_p, [C.c_int]),
    "sbfe_num_frames": (_I64, [_I64, _I32]),
    "sbfe_fbank_workspace_bytes": (_I64, [_DESC, _I64, _I64]),
    "sbfe_fbank_forward": (C.c_int, [_DESC, _P, _I64, _I64, _I64, _P, _P, _P, _F, _P, _I64, _P]),
    "sbfe_fbank_forward_ragged": (C.c_int, [_DESC, _P, _P, _P, _I64, _I64, _P, _P, _P, _F, _P]),
    "sbfe_fbank_forward_ragged_pcm16": (C.c_int, [_DESC, _P, _P, _P, _I64, _I64, _F, _P, _P, _P, _F,
                                                  _P]),
    "sbfe_fbank_tile_frames": (_I32, [_DESC]),
    "sbfe_fbank_ragged_supported": (_I32, [_DESC]),
    "sbfe_fbank_tile_stats_floats": (_I32, [_DESC]),
    "sbfe_norm_stats_tiles": (C.c_int, [_P, _P, _I64, _I64, _I32, _I32, _I32, _F, _P, _I32, _F, _P,
                                        _P, _P]),
    "sbfe_norm_sentence_tiles": (C.c_int, [_P, _P, _I64, _I64, _I32, _I32, _I32, _F, _P, _I32, _F,
                                           _I32, _P, _P]),
    "sbfe_stft_frames": (_I64, [_DESC, _I64, _I32]),
    "sbfe_stft_forward": (C.c_int, [_DESC, _P, _I64, _I64, _I64, _P, _I32, _I32, _P]),
    "sbfe_istft_workspace_bytes": (_I64, [_I64, _I64, _I32]),
    "sbfe_istft_forward": (C.c_int, [_P, _I64, _I64, _I32, _I32, _P, _I32, _I64, _P, _P, _I64, _P]),
    "sbfe_spectral_magnitude": (C.c_int, [_P, _I64, _F, _I32, _F, _P, _P]),
    "sbfe_filterbank_forward": (C.c_int, [_DESC, _P, _I64, _I64, _P, _P, _P]),
    "sbfe_post_forward": (C.c_int, [_P, _I64, _I64, _I32, _P, _F, _P, _I32, _I32, _I32, _I32,
                                    _I32, _I32, _P, _P]),
    "sbfe_norm_workspace_bytes": (_I64, [_I64, _I64, _I32]),
    "sbfe_norm_stats": (C.c_int, [_P, _P, _I64, _I64, _I32, _I32, _I32, _F, _P, _F, _P, _P, _P,
                                  _I64, _P]),
    "sbfe_norm_stats_static": (C.c_int, [_P, _P, _I64, _I64, _I32, _I32, _I32, _F, _P, _F, _P, _P, _P]),
    "sbfe_norm_reduce": (C.c_int, [_P, _P, _I64, _I32, _P, _P]),
    "sbfe_norm_update": (C.c_int, [_P, _I32, _I32, _F, _P, _P, _P]),
    "sbfe_norm_apply": (C.c_int, [_P, _I64, _I64, _I32, _P, _P, _I32, _P, _F, _I32, _P, _P]),
    "sbfe_norm_speaker_update": (C.c_int, [_P, _P, _I64, _I32, _P, _P, _P, _P, _P, _P, _P, _P]),
    "sbfe_specaugment_forward": (C.c_int, [_P, _I64, _I64, _I32, _I32, _I32, _I32, _P, _P, _I32, _P,
                                           _P, _I32, _P, _F, _P, _P]),
    "sbfe_wav_drop_chunks": (C.c_int, [_P, _I64, _I64, _I32, _P, _I32, _I64, _P, _P, _P]),
    "sbfe_wav_fir": (C.c_int, [_P, _I64, _I64, _P, _I32, _I32, _P, _P]),
    "sbfe_wav_clip": (C.c_int, [_P, _I64, _F, _P, _P]),
    "sbfe_wav_notch_taps": (C.c_int, [_P, _I32, _F, _I32, _P, _P]),
    "sbfe_wav_abs_chunks": (_I64, [_I64]),
    "sbfe_wav_abs_sums": (C.c_int, [_P, _I64, _I64, _P, _P]),
    "sbfe_wav_add_noise": (C.c_int, [_P, _P, _I64, _I64, _P, _P, _P, _P, _P, _P]),
    "sbfe_wav_peak_normalize": (C.c_int, [_P, _I64, _I64, _P, _P]),
    "sbfe_wav_add_babble": (C.c_int, [_P, _I64, _I64, _I32, _P, _P, _P, _P, _P]),
    "sbfe_wav_resample": (C.c_int, [_P, _I64, _I64, _P, _P, _I32, _I32, _I32, _I64, _P, _P]),
    "sbfe_mean_workspace_bytes": (_I64, []),
    "sbfe_mean": (C.c_int, [_P, _I64, _P, _P, _I64, _P]),
}

ABI_VERSION = 14
_lock = threading.Lock()
_lib = None


def loaded_path():
    return _build.LIB


def lib():
    """Load (building first when the sources are newer and nvcc exists)."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        try:
            if _build.is_stale():
                _build.build()
        except Exception as e:  # no nvcc on this machine
            if not os.path.exists(_build.LIB):
                raise SbfeUnavailable(
                    "libsbfe.so is not built and cannot be built here (%s). "
                    "Run `python -m speechbrain_b200._build`; there is no CPU fallback." % e
                ) from e
        try:
            handle = C.CDLL(_build.LIB)
        except OSError as e:
            raise SbfeUnavailable("cannot load %s: %s" % (_build.LIB, e)) from e
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(handle, name)
            fn.restype, fn.argtypes = res, args
        if handle.sbfe_abi_version() != ABI_VERSION:
            raise SbfeUnavailable("libsbfe.so ABI version mismatch")
        _lib = handle
        return _lib


def check(rc):
    if rc != 0:
        raise RuntimeError("sbfe call failed (%d): %s" % (rc, lib().sbfe_error_string(rc).decode()))
