"""speechbrain_b200: B200-native (sm_100a) acoustic feature frontend for SpeechBrain.

Drop-in replacements for ``speechbrain.lobes.features.{Fbank, MFCC}`` and
``speechbrain.processing.features.{STFT, spectral_magnitude, Filterbank, DCT,
Deltas, ContextWindow, InputNormalization}`` (and, next to the path,
``speechbrain.lobes.augment.SpecAugment``) backed by hand-written CUDA kernels
behind a C ABI (include/sbfe.h).  See DESIGN.md / INTEGRATION.md.
"""
from .dataio import RaggedBatch  # noqa: F401
from .lobes.augment import SpecAugment, TimeDomainSpecAugment  # noqa: F401
from .lobes.features import MFCC, Fbank, pair  # noqa: F401
from .processing.features import (  # noqa: F401
    DCT, ISTFT, STFT, ContextWindow, Deltas, Filterbank, InputNormalization, spectral_magnitude)
from .processing.speech_augmentation import (  # noqa: F401
    AddBabble, AddNoise, DoClip, DropChunk, DropFreq, Resample, SpeedPerturb)

__version__ = "0.1.0"

_PROCESSING = ("STFT", "ISTFT", "spectral_magnitude", "Filterbank", "DCT", "Deltas", "ContextWindow",
               "InputNormalization")
_LOBES = ("Fbank", "MFCC")


def install():
    """Rebind the frontend names inside an importable ``speechbrain`` so unchanged
    recipes / YAML (``!new:speechbrain.lobes.features.Fbank``) construct the B200
    classes.  Opt-in; call before ``load_hyperpyyaml``.  Returns the names patched."""
    import importlib

    from .processing import normalization

    patched = []
    proc = importlib.import_module("speechbrain.processing.features")
    lobes = importlib.import_module("speechbrain.lobes.features")
    here = importlib.import_module(__name__)
    for name in _PROCESSING:
        setattr(proc, name, getattr(here, name))
        patched.append("speechbrain.processing.features." + name)
    for name in _PROCESSING:
        if hasattr(lobes, name):  # lobes imported them by name
            setattr(lobes, name, getattr(here, name))
    for name in _LOBES:
        setattr(lobes, name, getattr(here, name))
        patched.append("speechbrain.lobes.features." + name)
    try:  # the step after the path (lobes/augment.py pulls in torchaudio; optional)
        aug = importlib.import_module("speechbrain.lobes.augment")
        aug.SpecAugment = here.SpecAugment
        aug.TimeDomainSpecAugment = here.TimeDomainSpecAugment
        patched += ["speechbrain.lobes.augment.SpecAugment",
                    "speechbrain.lobes.augment.TimeDomainSpecAugment"]
        sa = importlib.import_module("speechbrain.processing.speech_augmentation")
        for name in ("DropChunk", "DropFreq", "DoClip", "Resample", "SpeedPerturb", "AddBabble"):
            setattr(sa, name, getattr(here, name))
            patched.append("speechbrain.processing.speech_augmentation." + name)
    except Exception:
        pass
    normalization.register_checkpoint_hooks_if_available()
    return patched
