"""Builds libpmtech_b200.so in-tree with nvcc for sm_100a (no torch involved).

    python -m pmtech_b200.build [--force]

The shared object lands in pmtech_b200/lib/ (git-ignored; it travels to the GPU box with the
gpurun snapshot).  -fmad=false and -ffp-contract=off pin "no FMA contraction" on both the
device and the host side of the library (the device code also uses explicit __f*_rn intrinsics).
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIBDIR = os.path.join(HERE, "lib")
SO = os.path.join(LIBDIR, "libpmtech_b200.so")
SOURCES = ["pm_kernels.cu", "pm_capi.cu", "pm_exchange.cu", "pm_noalloc.cu", "pm_camera.cpp", "pm_pms.cpp"]
HEADERS = ["pm_kernels.h", "pm_math.cuh", "pm_camera_math.h", os.path.join("..", "..", "include", "pmtech_b200.h")]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17", "-fmad=false",
    "-Xcompiler", "-fPIC,-ffp-contract=off,-fno-fast-math",
    "-cudart", "static",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found; the CUDA path cannot be built")


def _stale(target: str, deps) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(LIBDIR, exist_ok=True)
    srcs = [os.path.join(CSRC, s) for s in SOURCES]
    import glob
    deps = (srcs + [os.path.normpath(os.path.join(CSRC, h)) for h in HEADERS] + glob.glob(os.path.join(CSRC, "*.h"))
            + glob.glob(os.path.join(CSRC, "*.cuh")) + [os.path.abspath(__file__)])
    if force or _stale(SO, deps):
        cmd = [_nvcc()] + NVCC_FLAGS + ["-shared", "-o", SO] + srcs
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
            print(" ".join(cmd))
        subprocess.check_call(cmd)
    return SO


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
