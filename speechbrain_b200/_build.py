"""Builds speechbrain_b200/csrc/libsbfe.so in-tree with nvcc for sm_100a.

    python -m speechbrain_b200._build [--force]

The library is a plain C-ABI shared object (include/sbfe.h): no torch, no
pybind.  It is git-ignored but travels to the GPU box with the repo snapshot.
"""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(CSRC, "libsbfe.so")
SOURCES = ["api.cu", "fbank400.cu", "fbank400_v3.cu", "fbank512.cu", "generic.cu", "post.cu", "norm.cu", "augment.cu", "istft.cu", "timedomain.cu", "mel_umma.cu"]
HEADERS = ["common.cuh", "fft_codelets.cuh", "fft_packed.cuh", os.path.join("..", "..", "include", "sbfe.h")]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "--shared", "-Xcompiler", "-fPIC",
    "-Xcompiler", "-fvisibility=hidden",
    "--cudart", "shared",
]


def _nvcc():
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: cannot build libsbfe.so")


def is_stale():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(os.path.join(CSRC, f)) > t for f in SOURCES + HEADERS)


def build(force=False, verbose=False):
    if not force and not is_stale():
        return LIB
    extra = os.environ.get("SBFE_NVCC_EXTRA", "").split()  # experiment switches (profiles/kbench.py)
    cmd = [_nvcc()] + NVCC_FLAGS + extra + (["-Xptxas", "-v"] if verbose else []) + SOURCES + ["-o", LIB]
    res = subprocess.run(cmd, cwd=CSRC, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + res.stdout + res.stderr)
    if verbose:
        print(res.stderr)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
