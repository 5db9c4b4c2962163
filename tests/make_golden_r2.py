"""Round-2 fixtures from the UNMODIFIED reference (see make_golden.py for the loading trick):

    python tests/make_golden_r2.py

* tests/golden/frontend_golden_r2.npz: the CASES_R2 of golden_cases.py (filterbank jitter in
  training mode, n_fft = 400 with other hops, n_fft = 512 at tile-spanning sizes, gaussian
  filters, normalized STFT) and the eval-step outputs of the two checkpoints below;
* tests/golden/ref_normalizer_{global,speaker}.ckpt: ``normalizer.ckpt`` files written by the
  reference's own ``InputNormalization._save`` (processing/features.py:1205-1215);
* and it checks the REVERSE direction on the spot: a file written by the product's ``_save``
  is loaded by the reference class and must give the same eval output.
"""
import os
import sys
import tempfile
import warnings

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import golden_cases as gc  # noqa: E402
from make_golden import load_reference, run_case  # noqa: E402


def run_case_r2(name, spec, lobes, proc):
    if "train_seed" not in spec:
        return run_case(name, spec, lobes, proc)
    x = spec["make_input"]()
    cls = lobes.MFCC if spec["kind"] == "mfcc" else lobes.Fbank
    mod = cls(**spec["kwargs"]).train()
    torch.manual_seed(spec["train_seed"])
    feats = mod(torch.from_numpy(x))
    torch.manual_seed(spec["train_seed"])
    out = {"in_ck": gc.input_checksum(x), "ref32": feats.numpy().copy(),
           "rand": torch.rand(2).numpy().copy(),
           "fc": mod.compute_fbanks.f_central.numpy().copy(),
           "band": mod.compute_fbanks.band.numpy().copy(),
           "window": mod.compute_STFT.window.numpy().copy()}
    return {f"{name}/{k}": v for k, v in out.items()}


def build_state(norm_cls, norm_type):
    norm = norm_cls(norm_type=norm_type).train()
    for st in gc.CKPT_STEPS:
        x = torch.from_numpy(gc.synth_feats(st["seed"], gc.CKPT_SHAPE))
        norm(x, torch.tensor(st["lengths"]), spk_ids=torch.tensor(st["spk_ids"]).unsqueeze(1),
             epoch=st["epoch"])
    return norm


def eval_step(norm):
    norm.eval()
    x = torch.from_numpy(gc.synth_feats(gc.CKPT_EVAL["seed"], gc.CKPT_SHAPE))
    y = norm(x, torch.tensor(gc.CKPT_EVAL["lengths"]),
             spk_ids=torch.tensor(gc.CKPT_EVAL["spk_ids"]).unsqueeze(1))
    return y.numpy().copy()


def main():
    warnings.filterwarnings("ignore")
    torch.manual_seed(0)
    torch.set_num_threads(1)
    lobes, proc = load_reference()
    blob = {}
    for name, spec in gc.CASES_R2.items():
        blob.update(run_case_r2(name, spec, lobes, proc))
        print("ok", name)
    import speechbrain_b200 as sb

    for norm_type, path in (("global", gc.REF_CKPT_GLOBAL), ("speaker", gc.REF_CKPT_SPEAKER)):
        ref = build_state(proc.InputNormalization, norm_type)
        ref._save(path)
        blob[f"ckpt_{norm_type}/eval"] = eval_step(ref)
        blob[f"ckpt_{norm_type}/count"] = np.array(ref.count)
        # reverse direction: product loads the reference's file, saves its own, the reference
        # class loads THAT and must reproduce the eval output (host-only: no kernel involved)
        prod = sb.InputNormalization(norm_type=norm_type)
        prod._load(path, device="cpu")
        with tempfile.TemporaryDirectory() as d:
            p2 = os.path.join(d, "normalizer.ckpt")
            prod._save(p2)
            back = proc.InputNormalization(norm_type=norm_type)
            # the reference's _load_statistics_dict reads self.device_inp (:1180), which its
            # __init__ never sets: with a non-empty speaker dictionary the unmodified class
            # cannot load ANY checkpoint, its own included.  Supplied here, the class is
            # otherwise untouched.
            back.device_inp = torch.device("cpu")
            back._load(p2)
            assert np.array_equal(eval_step(back), blob[f"ckpt_{norm_type}/eval"]), norm_type
        print("ok ckpt", norm_type, os.path.getsize(path), "bytes")
    blob["meta/torch"] = np.array(torch.__version__)
    np.savez_compressed(gc.GOLDEN_R2_NPZ, **blob)
    print("wrote", gc.GOLDEN_R2_NPZ, os.path.getsize(gc.GOLDEN_R2_NPZ), "bytes")


if __name__ == "__main__":
    main()
