"""Tile programs for tests/test_tir_splice.py, traced by a REAL TileLang (`import tilelang` = the compiler, not this
repo's facade).  Same program shapes as tests/tile_programs.py, restricted to the keyword arguments that both the
reference snapshot (0.1.7) and the wheel of the image (0.1.9) accept (`wg_wait` left the 0.1.9 `T.gemm` signature)."""
import tilelang  # noqa: F401
import tilelang.language as T


def tmem_accumulator(M, N, K, bM=128, bN=256, bK=128, dt="bfloat16", out_dt="bfloat16", stages=2, threads=256):
    # the program shape of examples/gemm_sm100/gemm_tcgen5mma.py: tensor-memory accumulator, user mbarrier
    @T.prim_func
    def main(A: T.Tensor((M, K), dt), B: T.Tensor((N, K), dt), C: T.Tensor((M, N), out_dt)):
        with T.Kernel(T.ceildiv(N, bN), T.ceildiv(M, bM), threads=threads) as (bx, by):
            A_s = T.alloc_shared((bM, bK), dt)
            B_s = T.alloc_shared((bN, bK), dt)
            C_t = T.alloc_tmem([bM, bN], "float")
            mbar = T.alloc_barrier(1)
            C_l = T.alloc_fragment((bM, bN), "float")
            C_s = T.alloc_shared((bM, bN), out_dt)
            for k in T.Pipelined(T.ceildiv(K, bK), num_stages=stages):
                T.copy(A[by * bM, k * bK], A_s)
                T.copy(B[bx * bN, k * bK], B_s)
                T.gemm(A_s, B_s, C_t, False, True, mbar=mbar, clear_accum=k == 0)
                T.mbarrier_wait_parity(mbar, k % 2)
            T.copy(C_t, C_l)
            T.copy(C_l, C_s)
            T.copy(C_s, C[by * bM, bx * bN])

    return main


def nn_relu(M, N, K, bM=128, bN=128, bK=32, dt="float16"):
    # the program shape of examples/quickstart.py: B is [K, N], relu fused after the K loop
    @T.prim_func
    def main(A: T.Tensor((M, K), dt), B: T.Tensor((K, N), dt), C: T.Tensor((M, N), dt)):
        with T.Kernel(T.ceildiv(N, bN), T.ceildiv(M, bM), threads=128) as (bx, by):
            A_s = T.alloc_shared((bM, bK), dt)
            B_s = T.alloc_shared((bK, bN), dt)
            C_l = T.alloc_fragment((bM, bN), "float")
            T.use_swizzle(panel_size=10, enable=True)
            T.clear(C_l)
            for ko in T.Pipelined(T.ceildiv(K, bK), num_stages=3):
                T.copy(A[by * bM, ko * bK], A_s)
                T.copy(B[ko * bK, bx * bN], B_s)
                T.gemm(A_s, B_s, C_l)
            for i, j in T.Parallel(bM, bN):
                C_l[i, j] = T.max(C_l[i, j], 0)
            T.copy(C_l, C[by * bM, bx * bN])

    return main


def nt_fragment(M, N, K, bM=128, bN=128, bK=64, dt="float16", out_dt=None, stages=3):
    # the program shape of benchmark/matmul and benchmark/matmul_fp8: register accumulator, NT layout
    out_dt = out_dt or dt

    @T.prim_func
    def main(A: T.Tensor((M, K), dt), B: T.Tensor((N, K), dt), C: T.Tensor((M, N), out_dt)):
        with T.Kernel(T.ceildiv(N, bN), T.ceildiv(M, bM), threads=128) as (bx, by):
            A_s = T.alloc_shared((bM, bK), dt)
            B_s = T.alloc_shared((bN, bK), dt)
            C_l = T.alloc_fragment((bM, bN), "float")
            C_s = T.alloc_shared((bM, bN), out_dt)
            T.clear(C_l)
            for k in T.Pipelined(T.ceildiv(K, bK), num_stages=stages):
                T.copy(A[by * bM, k * bK], A_s)
                T.copy(B[bx * bN, k * bK], B_s)
                T.gemm(A_s, B_s, C_l, transpose_B=True)
            T.copy(C_l, C_s)
            T.copy(C_s, C[by * bM, bx * bN])

    return main
