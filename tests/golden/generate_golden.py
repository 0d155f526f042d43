"""Generates the golden fixtures under tests/golden/ from the reference LINEAGE's own code.

The reference snapshot (/root/reference, TileLang 0.1.7 + mesh) cannot be built or imported
(3rdparty/tvm is absent).  The same lineage's later release
(tilelang 0.1.9, pip-installed in the build container) CAN be imported, so this script uses it:

 1. tcgen05_lowering.json — lowers the reference's sm_100 GEMM program
    (examples/gemm_sm100/gemm_tcgen5mma.py shape: tmem accumulator, TMA/WS disabled) for
    `cuda -arch=sm_100a` over a grid of tile shapes / dtypes and records, from the GENERATED
    CUDA source, the tcgen05 instruction descriptor, the shared-memory descriptor arguments
    (LBO, SBO, base offset, lbo mode, swizzle), the per-MMA K-advance byte offsets and the TMEM
    column count.  These pin oracle/descriptors.py (and through it tlb200_instr_desc /
    tlb200_smem_desc / tlb200_tcgen5_meta).
 2. cpu_c_target_matmul.npz — compiles and RUNS the reference's CPU-target scalar-loop matmul
    (testing/python/cpu/test_tilelang_cpu_gemm.py:66-98) on seeded fp16 inputs and stores
    inputs + output.  This pins the numerical oracle (oracle/gemm_oracle.c: fp32 accumulate in
    k order, one rounding to the output type) bit-for-bit.

Run in the build container only:  python tests/golden/generate_golden.py
Nothing here is needed at test time; the fixtures are committed.
"""
import json
import os
import re

import numpy as np
import torch

import tilelang
import tilelang.language as T
from tilelang import tvm

HERE = os.path.dirname(os.path.abspath(__file__))


def sm100_program(M, N, K, bM, bN, bK, in_dtype, out_dtype, accum="float", stages=2, threads=256):

    @T.prim_func
    def main(A: T.Tensor((M, K), in_dtype), B: T.Tensor((N, K), in_dtype), C: T.Tensor((M, N), out_dtype)):
        with T.Kernel(T.ceildiv(N, bN), T.ceildiv(M, bM), threads=threads) as (bx, by):
            A_s = T.alloc_shared((bM, bK), in_dtype)
            B_s = T.alloc_shared((bN, bK), in_dtype)
            C_t = T.alloc_tmem([bM, bN], accum)
            mbar = T.alloc_barrier(1)
            C_l = T.alloc_fragment((bM, bN), accum)
            C_s = T.alloc_shared((bM, bN), out_dtype)
            for k in T.Pipelined(T.ceildiv(K, bK), num_stages=stages):
                T.copy(A[by * bM, k * bK], A_s)
                T.copy(B[bx * bN, k * bK], B_s)
                T.gemm(A_s, B_s, C_t, False, True, mbar=mbar, clear_accum=k == 0)
                T.mbarrier_wait_parity(mbar, k % 2)
            T.copy(C_t, C_l)
            T.copy(C_l, C_s)
            T.copy(C_s, C[by * bM, bx * bN])

    return main


def lower_sm100(bM, bN, bK, in_dtype, out_dtype):
    tgt = tvm.target.Target("cuda -arch=sm_100a")
    cfg = {
        tilelang.PassConfigKey.TL_DISABLE_TMA_LOWER: True,
        tilelang.PassConfigKey.TL_DISABLE_WARP_SPECIALIZED: True,
    }
    M, N, K = 2 * bM, 2 * bN, 2 * bK
    with tvm.transform.PassContext(opt_level=3, config=cfg), tgt:
        return tilelang.lower(sm100_program(M, N, K, bM, bN, bK, in_dtype, out_dtype), target=tgt).kernel_source


def parse(src):
    rec = {}
    m = re.search(r"tcgen05mma_(ws_)?ss<tl::DataType::(\w+), (\w+)>\((.*?)\);", src)
    if not m:
        return None
    rec["ws"] = bool(m.group(1))
    rec["kind"] = m.group(2)
    call = m.group(4)
    idesc = re.search(r"static_cast<uint32_t>\((\d+)\)", call)
    rec["instr_desc"] = int(idesc.group(1))
    offs = re.findall(r"uint64_t\(desc_[ab](?:_\d+)? \+ (.*?)\)\)?,", call)
    rec["k_offset_exprs"] = offs[:2]
    init = re.findall(r"initialize_tcgen05_descriptor\(desc_([ab]), .*?\]\)\), (\d+), (\d+), (\d+), (\d+), (\d+)\);", src)
    for which, lbo, sbo, base, lbo_mode, swz in init[:2]:
        rec["desc_" + which] = dict(lbo16=int(lbo), sbo16=int(sbo), base_offset=int(base), lbo_mode=int(lbo_mode),
                                    swizzle=int(swz))
    loop = re.search(r"for \(int ki = 0; ki < (\d+); \+\+ki\)", src)
    rec["k_atoms"] = int(loop.group(1)) if loop else 1
    t = re.search(r"tmem_allocate\(.*?, (\d+)\)", src)
    rec["tmem_cols"] = int(t.group(1))
    return rec


def gen_lowering():
    out = []
    grid = []
    for bM in (128, 64):
        for bN in (64, 128, 256):
            for bK in (64, 128):
                grid.append((bM, bN, bK, "bfloat16", "bfloat16"))
    grid += [(128, 256, 128, "float16", "float16"), (128, 128, 64, "float16", "float16"),
             (128, 256, 128, "float8_e4m3", "bfloat16"), (128, 128, 128, "float8_e4m3", "bfloat16"),
             (64, 256, 128, "float8_e4m3", "bfloat16"), (128, 256, 128, "float8_e5m2", "bfloat16"),
             (128, 48, 64, "bfloat16", "bfloat16"), (128, 80, 64, "bfloat16", "bfloat16")]
    for bM, bN, bK, idt, odt in grid:
        err = None
        try:
            src = lower_sm100(bM, bN, bK, idt, odt)
            rec = parse(src)
        except Exception as e:  # noqa: BLE001 - record which shapes the lineage rejects
            rec = None
            err = f"{type(e).__name__}: {str(e).splitlines()[0][:160]}"
        entry = dict(block_M=bM, block_N=bN, block_K=bK, in_dtype=idt, out_dtype=odt)
        if rec:
            entry.update(rec)
        else:
            entry["unsupported"] = err or "no tcgen05 call in source"
        out.append(entry)
        print(entry)
    with open(os.path.join(HERE, "tcgen05_lowering.json"), "w") as f:
        json.dump(dict(generator="tilelang " + tilelang.__version__, target="cuda -arch=sm_100a", records=out), f,
                  indent=1)


def cpu_matmul_program(M, N, K, bM, bN, bK, dtype="float16", accum_dtype="float"):
    # testing/python/cpu/test_tilelang_cpu_gemm.py:66-98, verbatim structure.
    @T.prim_func
    def matmul(A: T.Tensor((M, K), dtype), B: T.Tensor((K, N), dtype), C: T.Tensor((M, N), dtype)):
        with T.Kernel(T.ceildiv(N, bN), T.ceildiv(M, bM), is_cpu=True) as (bx, by):
            A_local = T.alloc_local((bM, bK), dtype)
            B_local = T.alloc_local((bK, bN), dtype)
            C_local = T.alloc_local((bM, bN), accum_dtype)
            for p in T.serial(bM):
                for w in T.serial(bN):
                    C_local[p, w] = 0
            for ko in T.serial(K // bK):
                for i in T.serial(bM):
                    for k in T.serial(bK):
                        A_local[i, k] = A[by * bM + i, ko * bK + k]
                for k in T.serial(bK):
                    for j in T.serial(bN):
                        B_local[k, j] = B[ko * bK + k, bx * bN + j]
                for i in T.serial(bM):
                    for j in T.serial(bN):
                        for k in T.serial(bK):
                            C_local[i, j] += A_local[i, k] * B_local[k, j]
            for i in T.serial(bM):
                for j in T.serial(bN):
                    C[by * bM + i, bx * bN + j] = C_local[i, j]

    return matmul


def gen_cpu():
    M, N, K = 96, 80, 64
    func = cpu_matmul_program(M, N, K, 32, 16, 16)
    with tvm.target.Target("c"):
        kern = tilelang.compile(func, out_idx=-1, execution_backend="cython")
    torch.manual_seed(0)
    A = torch.randn(M, K, dtype=torch.float16)
    B = torch.randn(K, N, dtype=torch.float16)
    C = kern(A, B)
    np.savez_compressed(os.path.join(HERE, "cpu_c_target_matmul.npz"), A=A.view(torch.int16).numpy(),
                        B=B.view(torch.int16).numpy(), C=C.view(torch.int16).numpy(),
                        meta=np.array([M, N, K], dtype=np.int32))
    ref = (A.float() @ B.float()).half()
    print("cpu golden: max |C - torch| =", (C.float() - ref.float()).abs().max().item(), "bit-equal:",
          bool((C.view(torch.int16) == ref.view(torch.int16)).all()))


if __name__ == "__main__":
    gen_lowering()
    gen_cpu()
