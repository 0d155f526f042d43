"""In-tree build of libtlb200.so with nvcc for sm_100a (cross-compiles without a GPU).

The reference compiles every JIT kernel with `nvcc --shared -gencode arch=compute_100a,code=sm_100a`
(tilelang/jit/adapter/libgen.py:53-158, tilelang/contrib/nvcc.py:470-475); this backend is
ahead-of-time: one library, built once.
"""
from __future__ import annotations

import os
import shutil
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(_HERE)
CSRC = os.path.join(_HERE, "csrc")
LIB_DIR = os.path.join(_HERE, "lib")
LIB = os.path.join(LIB_DIR, "libtlb200.so")
SOURCES = [os.path.join(CSRC, "tlb200.cu")]
DEPS = SOURCES + [os.path.join(CSRC, f) for f in ("gemm_sm100.cuh", "gemm_blockscaled_sm100.cuh", "ptx.cuh")] + [
    os.path.join(ROOT, "include", "tlb200.h")]


def nvcc_path() -> str:
    p = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(p):
        raise RuntimeError("nvcc not found: libtlb200.so needs the CUDA 12.9 toolkit to build")
    return p


def is_stale() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(d) > t for d in DEPS if os.path.exists(d))


def build_library(force: bool = False, verbose: bool = False) -> str:
    if not force and not is_stale():
        return LIB
    os.makedirs(LIB_DIR, exist_ok=True)
    cmd = [nvcc_path(), "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
           "-shared", "-Xcompiler", "-fPIC", "-I", os.path.join(ROOT, "include"), "-I", CSRC, "-o", LIB] + SOURCES
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + res.stdout + res.stderr)
    if verbose:
        print(res.stderr)
    return LIB


if __name__ == "__main__":
    print(build_library(force=True, verbose=True))
