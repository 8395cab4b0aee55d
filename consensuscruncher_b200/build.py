"""Builds libccgpu.so (the sm_100a CUDA library behind the C ABI of include/ccgpu.h)
in-tree with nvcc.  No GPU is needed to build; one is needed to load and run."""
from __future__ import annotations

import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libccgpu.so")
SOURCES = ["ccgpu.cu"]
HEADERS = ["kernels.cuh", "radix_onesweep.cuh", "scan.cuh", "vote.cuh", "group_hash.cuh", "group_tile.cuh", "group_pair.cuh", os.path.join("..", "..", "include", "ccgpu.h")]

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-shared", "-Xcompiler", "-fPIC", "-cudart", "static"]


def nvcc_path():
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def stale():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(os.path.join(CSRC, f)) > t for f in SOURCES + HEADERS)


def build_lib(force=False, verbose=False):
    """Compile csrc/*.cu into consensuscruncher_b200/libccgpu.so if it is out of date."""
    if not force and not stale():
        return LIB
    cmd = [nvcc_path()] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + \
          ["-o", LIB] + [os.path.join(CSRC, s) for s in SOURCES]
    out = subprocess.run(cmd, capture_output=True, text=True)
    if out.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + out.stdout + out.stderr)
    if verbose:
        print(out.stderr)
    return LIB


PACK_LIB = os.path.join(HERE, "libccpack.so")
PACK_SRCS = [os.path.join(CSRC, f) for f in ("ccpack.cpp", "ccemit.cpp")]
PACK_DEPS = PACK_SRCS + [os.path.join(CSRC, "ccpack.hpp")]


def build_pack_lib(force=False):
    """Compile csrc/ccpack.cpp + ccemit.cpp (host-only: BGZF inflate / deflate, packing, record
    assembly) into libccpack.so."""
    if not force and os.path.exists(PACK_LIB) and all(os.path.getmtime(PACK_LIB) >= os.path.getmtime(f) for f in PACK_DEPS):
        return PACK_LIB
    cxx = shutil.which("g++") or "g++"
    cmd = [cxx, "-O2", "-std=c++17", "-shared", "-fPIC", "-o", PACK_LIB] + PACK_SRCS + ["-lz", "-lpthread"]
    out = subprocess.run(cmd, capture_output=True, text=True)
    if out.returncode != 0:
        raise RuntimeError("g++ failed:\n" + out.stdout + out.stderr)
    return PACK_LIB


if __name__ == "__main__":
    print(build_lib(force=True, verbose=True))
    print(build_pack_lib(force=True))
