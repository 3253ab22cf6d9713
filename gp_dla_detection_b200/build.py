"""Build libdla_b200.so in-tree with nvcc for sm_100a (no torch, no JIT cache).

    python -m gp_dla_detection_b200.build            # build if stale
    python -m gp_dla_detection_b200.build --force
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG_DIR, "csrc")
LIB_PATH = os.path.join(PKG_DIR, "libdla_b200.so")
SOURCES = ["dla_b200.cu"]
HEADERS = ["kernels.cuh", "slices.cuh", "slices_path.cuh", "faddeeva.cuh", "faddeeva_table.cuh", "voigt_constants.cuh",
           os.path.join("..", "..", "include", "dla_b200.h")]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "--shared", "-Xcompiler", "-fPIC",
    "-cudart", "static", "-ldl",
    "-Xptxas", "-v",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: libdla_b200.so cannot be built (there is no CPU fallback)")


def is_stale() -> bool:
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    for f in SOURCES + HEADERS:
        p = os.path.join(CSRC, f)
        if os.path.exists(p) and os.path.getmtime(p) > t:
            return True
    return False


def build_library(force: bool = False, verbose: bool = False) -> str:
    if not force and not is_stale():
        return LIB_PATH
    extra = os.environ.get("DLA_NVCC_EXTRA", "").split()   # tuning builds only (e.g. -DDLA_K2P_KC=16 -DDLA_K2P_STAGES=2)
    cmd = [_nvcc(), *NVCC_FLAGS, *extra, "-o", LIB_PATH, *[os.path.join(CSRC, s) for s in SOURCES]]
    env = dict(os.environ)
    env.pop("CC", None)  # the image's CC=/opt/gcc/bin/gcc is not a usable host compiler for nvcc
    env.pop("CXX", None)
    res = subprocess.run(cmd, capture_output=True, text=True, env=env)
    log = res.stdout + res.stderr
    with open(os.path.join(PKG_DIR, "build.log"), "w") as f:
        f.write(" ".join(cmd) + "\n" + log)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + log[-8000:])
    if verbose:
        print(log)
    return LIB_PATH


if __name__ == "__main__":
    print(build_library(force="--force" in sys.argv, verbose=True))
