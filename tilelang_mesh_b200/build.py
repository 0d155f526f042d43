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
# the host side plus one translation unit per group of kernel instantiations: compiled in parallel, linked once
SOURCES = [os.path.join(CSRC, f) for f in ("tlb200.cu", "kernels_cg1_ns1.cu", "kernels_cg1_ns2.cu", "kernels_cg2_ns1.cu",
                                           "kernels_cg2_ns2.cu", "kernels_mc.cu", "kernels_blockscaled.cu", "kernels_chain.cu")]
DEPS = SOURCES + [os.path.join(CSRC, f) for f in ("gemm_sm100.cuh", "gemm_blockscaled_sm100.cuh", "gemm_chain_sm100.cuh", "ptx.cuh", "kernel_table.h")] + [
    os.path.join(ROOT, "include", "tlb200.h")]
OBJ_DIR = os.path.join(LIB_DIR, "obj")


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


def build_library(force: bool = False, verbose: bool = False, defines=(), out: str = LIB) -> str:
    if not force and out == LIB and not is_stale():
        return LIB
    from concurrent.futures import ThreadPoolExecutor
    obj_dir = OBJ_DIR if out == LIB else out + ".obj"
    os.makedirs(LIB_DIR, exist_ok=True)
    os.makedirs(obj_dir, exist_ok=True)
    base = [nvcc_path(), "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
            "-Xcompiler", "-fPIC", "-I", os.path.join(ROOT, "include"), "-I", CSRC] + ["-D" + d for d in defines]
    if verbose:
        base.insert(1, "-Xptxas=-v")

    def compile_one(src):
        obj = os.path.join(obj_dir, os.path.basename(src)[:-3] + ".o")
        res = subprocess.run(base + ["-c", src, "-o", obj], capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError(f"nvcc failed on {src}:\n" + res.stdout + res.stderr)
        return obj, res.stderr

    with ThreadPoolExecutor(max_workers=min(len(SOURCES), os.cpu_count() or 1)) as pool:
        results = list(pool.map(compile_one, SOURCES))
    if verbose:
        for _, err in results:
            print(err)
    res = subprocess.run([nvcc_path(), "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", out] + [o for o, _ in results],
                         capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc link failed:\n" + res.stdout + res.stderr)
    return out


if __name__ == "__main__":
    import sys
    # python -m tilelang_mesh_b200.build [-DNAME ...] [-o out.so]: development variants (e.g. -DTLB200_TRACE) go elsewhere
    defs = [a[2:] for a in sys.argv[1:] if a.startswith("-D")]
    out = sys.argv[sys.argv.index("-o") + 1] if "-o" in sys.argv else LIB
    print(build_library(force=True, verbose="-v" in sys.argv, defines=defs, out=out))
