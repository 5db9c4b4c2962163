"""Generate ``tests/golden/*.npz`` by running the UNMODIFIED reference modules
(``/root/reference/speechbrain``) on CPU in the build container.

    python tests/make_golden.py

The reference cannot travel to the GPU box, so its outputs are committed as
fixtures.  ``import speechbrain`` itself fails here (hyperpyyaml is absent), so
package shells with the real ``__path__`` are registered first and the two
feature files are imported directly.
"""
import os
import sys
import types
import wave
import warnings

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import golden_cases as gc  # noqa: E402

REF_ROOT = os.environ.get("SB_REFERENCE", "/root/reference")


def load_reference():
    ref = os.path.join(REF_ROOT, "speechbrain")
    if "speechbrain" not in sys.modules:
        def shell(name, path):
            m = types.ModuleType(name)
            m.__path__ = [path]
            sys.modules[name] = m
        shell("speechbrain", ref)
        for sub in ("utils", "processing", "nnet", "lobes", "dataio"):
            shell("speechbrain." + sub, os.path.join(ref, sub))
    import speechbrain.lobes.features as lobes
    import speechbrain.processing.features as proc
    return lobes, proc


def make_speech_fixture():
    path = os.path.join(REF_ROOT, "tests/samples/single-mic/example1.wav")
    with wave.open(path, "rb") as w:
        assert w.getframerate() == 16000 and w.getsampwidth() == 2 and w.getnchannels() == 1
        pcm = np.frombuffer(w.readframes(w.getnframes()), dtype=np.int16)
    start = 8000
    np.savez_compressed(gc.SPEECH_NPZ, pcm=pcm[start:start + 24000])


def to_double(mod):
    """fp64 anchor: the reference's fp32-built constants promoted to double."""
    mod.compute_STFT.window = mod.compute_STFT.window.double()
    fb = mod.compute_fbanks
    fb.band, fb.f_central, fb.all_freqs_mat = fb.band.double(), fb.f_central.double(), fb.all_freqs_mat.double()
    if hasattr(mod, "compute_dct"):
        mod.compute_dct.dct_mat = mod.compute_dct.dct_mat.double()
    mod.compute_deltas.kernel = mod.compute_deltas.kernel.double()
    return mod


def run_case(name, spec, lobes, proc):
    out = {}
    kind = spec["kind"]
    t = torch.from_numpy
    if kind in ("fbank", "mfcc", "fbank+norm"):
        x = spec["make_input"]()
        out["in_ck"] = gc.input_checksum(x)
        cls = lobes.MFCC if kind == "mfcc" else lobes.Fbank
        mod = cls(**spec["kwargs"])
        feats = mod(t(x))
        out["ref32"] = feats.numpy().copy()
        # the reference's own float32 constants (last-ulp CPU-kernel dependent)
        out["fc"] = mod.compute_fbanks.f_central.numpy().copy()
        out["band"] = mod.compute_fbanks.band.numpy().copy()
        out["window"] = mod.compute_STFT.window.numpy().copy()
        kw = spec["kwargs"]
        # rectangular filters are cast with .float() in the reference: no fp64 run
        if kw.get("filter_shape") != "rectangular" and not (
            kw.get("context") or (kind == "mfcc" and kw.get("context", True))
        ):
            out["ref64"] = to_double(cls(**kw))(t(x).double()).numpy()
        if kind == "fbank+norm":
            norm = proc.InputNormalization(**spec["norm"])
            lens = torch.tensor(spec["lengths"], dtype=torch.float32)
            out["norm32"] = norm(feats.clone(), lens).numpy().copy()
            norm64 = proc.InputNormalization(**spec["norm"])
            out["norm64"] = norm64(t(out["ref64"]).clone(), lens).numpy()
    elif kind == "stft":
        x = spec["make_input"]()
        out["in_ck"] = gc.input_checksum(x)
        out["ref32"] = proc.STFT(**spec["kwargs"])(t(x)).numpy().copy()
    elif kind == "magnitude":
        x = spec["make_input"]()
        out["in_ck"] = gc.input_checksum(x)
        out["ref32"] = proc.spectral_magnitude(t(x), **spec["kwargs"]).numpy().copy()
    elif kind in ("filterbank", "dct", "deltas", "context"):
        x = spec["make_input"]()
        out["in_ck"] = gc.input_checksum(x)
        cls = {"filterbank": proc.Filterbank, "dct": proc.DCT, "deltas": proc.Deltas,
               "context": proc.ContextWindow}[kind]
        out["ref32"] = cls(**spec["kwargs"])(t(x)).numpy().copy()
    elif kind == "norm_seq":
        norm = proc.InputNormalization(**spec["norm"])
        for i, st in enumerate(spec["steps"]):
            x = gc.synth_feats(st["seed"], spec["shape"])
            norm.train(st["training"])
            args = [t(x.copy()), torch.tensor(st["lengths"], dtype=torch.float32)]
            kw = {"epoch": st["epoch"]}
            if st["spk_ids"] is not None:
                kw["spk_ids"] = torch.tensor(st["spk_ids"]).unsqueeze(1)
            y = norm(*args, **kw)
            out[f"step{i}"] = y.numpy().copy()
            if spec["norm"]["norm_type"] == "global":
                out[f"step{i}_glob_mean"] = np.asarray(norm.glob_mean, dtype=np.float32)
                out[f"step{i}_glob_std"] = np.asarray(norm.glob_std, dtype=np.float32)
                out[f"step{i}_count"] = np.array(norm.count)
    else:
        raise ValueError(kind)
    return {f"{name}/{k}": v for k, v in out.items()}


def main():
    warnings.filterwarnings("ignore")
    torch.manual_seed(0)
    torch.set_num_threads(1)  # deterministic reduction order for the fixtures
    os.makedirs(gc.GOLDEN_DIR, exist_ok=True)
    make_speech_fixture()
    lobes, proc = load_reference()
    blob = {}
    for name, spec in gc.CASES.items():
        blob.update(run_case(name, spec, lobes, proc))
        print("ok", name)
    blob["meta/torch"] = np.array(torch.__version__)
    np.savez_compressed(gc.GOLDEN_NPZ, **blob)
    print("wrote", gc.GOLDEN_NPZ, os.path.getsize(gc.GOLDEN_NPZ), "bytes")


if __name__ == "__main__":
    main()
