"""Tile programs written against the reference's public API (T.prim_func / T.Kernel / T.copy /
T.gemm / T.Pipelined), used by the facade tests.  They are this repo's own programs: the four
program shapes the reference ships for this path — register-accumulator NT (benchmark/matmul),
NN + relu (examples/quickstart.py), tensor-memory accumulator with user mbarrier
(examples/gemm_sm100), fp8 (benchmark/matmul_fp8) — re-expressed compactly."""
import tilelang
import tilelang.language as T


def nt_fragment(M, N, K, bM=128, bN=128, bK=32, dtype="float16", out_dtype=None, stages=3, threads=128, raster=True,
                accum="float"):
    out_dtype = out_dtype or dtype

    @T.prim_func
    def prog(A: T.Tensor((M, K), dtype), B: T.Tensor((N, K), dtype), C: T.Tensor((M, N), out_dtype)):
        with T.Kernel(T.ceildiv(N, bN), T.ceildiv(M, bM), threads=threads) as (bx, by):
            a_s = T.alloc_shared((bM, bK), dtype)
            b_s = T.alloc_shared((bN, bK), dtype)
            acc = T.alloc_fragment((bM, bN), accum)
            c_s = T.alloc_shared((bM, bN), out_dtype)
            T.use_swizzle(panel_size=10, enable=raster)
            T.annotate_layout({c_s: tilelang.layout.make_swizzled_layout(c_s)})
            T.clear(acc)
            for k in T.Pipelined(T.ceildiv(K, bK), num_stages=stages):
                T.copy(A[by * bM, k * bK], a_s)
                T.copy(B[bx * bN, k * bK], b_s)
                T.gemm(a_s, b_s, acc, transpose_B=True, policy=T.GemmWarpPolicy.Square)
            T.copy(acc, c_s)
            T.copy(c_s, C[by * bM, bx * bN])

    return prog


def nn_relu(M, N, K, bM=128, bN=128, bK=32, dtype="float16"):

    @T.prim_func
    def prog(A: T.Tensor((M, K), dtype), B: T.Tensor((K, N), dtype), C: T.Tensor((M, N), dtype)):
        with T.Kernel(T.ceildiv(N, bN), T.ceildiv(M, bM), threads=128) as (bx, by):
            a_s = T.alloc_shared((bM, bK), dtype)
            b_s = T.alloc_shared((bK, bN), dtype)
            acc = T.alloc_fragment((bM, bN), "float")
            T.clear(acc)
            for ko in T.Pipelined(T.ceildiv(K, bK), num_stages=3):
                T.copy(A[by * bM, ko * bK], a_s)
                T.copy(B[ko * bK, bx * bN], b_s)
                T.gemm(a_s, b_s, acc)
            for i, j in T.Parallel(bM, bN):
                acc[i, j] = T.max(acc[i, j], 0)
            T.copy(acc, C[by * bM, bx * bN])

    return prog


def tmem_accumulator(M, N, K, bM=128, bN=256, bK=128, trans_A=False, trans_B=True, in_dtype="bfloat16",
                     out_dtype="bfloat16", stages=2, threads=256, accum_dtype="float"):
    a_shape, b_shape = ((K, M) if trans_A else (M, K)), ((N, K) if trans_B else (K, N))
    a_tile, b_tile = ((bK, bM) if trans_A else (bM, bK)), ((bN, bK) if trans_B else (bK, bN))

    @T.prim_func
    def prog(A: T.Tensor(a_shape, in_dtype), B: T.Tensor(b_shape, in_dtype), C: T.Tensor((M, N), out_dtype)):
        with T.Kernel(T.ceildiv(N, bN), T.ceildiv(M, bM), threads=threads) as (bx, by):
            a_s = T.alloc_shared(a_tile, in_dtype)
            b_s = T.alloc_shared(b_tile, in_dtype)
            c_t = T.alloc_tmem([bM, bN], accum_dtype)
            bar = T.alloc_barrier(1)
            c_l = T.alloc_fragment((bM, bN), accum_dtype)
            c_s = T.alloc_shared((bM, bN), out_dtype)
            for k in T.Pipelined(T.ceildiv(K, bK), num_stages=stages):
                T.copy(A[(k * bK, by * bM) if trans_A else (by * bM, k * bK)], a_s)
                T.copy(B[(bx * bN, k * bK) if trans_B else (k * bK, bx * bN)], b_s)
                T.gemm(a_s, b_s, c_t, trans_A, trans_B, mbar=bar, wg_wait=-1, clear_accum=k == 0)
                T.mbarrier_wait_parity(bar, k % 2)
            T.copy(c_t, c_l)
            T.copy(c_l, c_s)
            T.copy(c_s, C[by * bM, bx * bN])

    return prog


def elementwise_staging(M, N, K, bM=128, bN=128, bK=32, dtype="float16"):
    """B tile filled element by element inside T.Parallel (the form of examples/gemm/example_gemm_schedule.py)."""

    @T.prim_func
    def prog(A: T.Tensor((M, K), dtype), B: T.Tensor((K, N), dtype), C: T.Tensor((M, N), dtype)):
        with T.Kernel(T.ceildiv(N, bN), T.ceildiv(M, bM), threads=128) as (bx, by):
            a_s = T.alloc_shared((bM, bK), dtype)
            b_s = T.alloc_shared((bK, bN), dtype)
            acc = T.alloc_fragment((bM, bN), "float")
            T.clear(acc)
            for ko in T.Pipelined(T.ceildiv(K, bK), num_stages=3):
                T.copy(A[by * bM, ko * bK], a_s)
                for kk, j in T.Parallel(bK, bN):
                    b_s[kk, j] = B[ko * bK + kk, bx * bN + j]
                T.gemm(a_s, b_s, acc)
            T.copy(acc, C[by * bM, bx * bN])

    return prog


def persistent_primitive(M, N, K, bM=128, bN=256, bK=64, dtype="bfloat16", waves=148):
    """Tile walk written with T.Persistent (the form of examples/gemm/example_gemm_persistent.py:92-110)."""

    @T.prim_func
    def prog(A: T.Tensor((M, K), dtype), B: T.Tensor((K, N), dtype), C: T.Tensor((M, N), dtype)):
        with T.Kernel(waves, threads=256) as (block_id):
            a_s = T.alloc_shared((bM, bK), dtype)
            b_s = T.alloc_shared((bK, bN), dtype)
            acc = T.alloc_fragment((bM, bN), "float")
            c_s = T.alloc_shared((bM, bN), dtype)
            for tx, ty in T.Persistent([T.ceildiv(M, bM), T.ceildiv(N, bN)], waves, block_id):
                T.clear(acc)
                for k in T.Pipelined(T.ceildiv(K, bK), num_stages=3):
                    T.copy(A[tx * bM, k * bK], a_s)
                    T.copy(B[k * bK, ty * bN], b_s)
                    T.gemm(a_s, b_s, acc)
                T.copy(acc, c_s)
                T.copy(c_s, C[tx * bM, ty * bN])

    return prog


def blockscaled_2xacc(M, N, K, bN=128, out_dtype="bfloat16", in_dtype="float8_e4m3", bM=128, bK=128, group=128,
                      swap_scale_operands=False):
    """fp8 contraction with per-(row, 128 K) and per-(128 N, 128 K) scales, promoted every K block
    (the form of examples/deepseek_deepgemm/example_deepgemm_fp8_2xAcc.py:43-83)."""

    @T.prim_func
    def prog(A: T.Tensor((M, K), in_dtype), B: T.Tensor((N, K), in_dtype), C: T.Tensor((M, N), out_dtype),
             sa: T.Tensor((M, T.ceildiv(K, group)), "float32"),
             sb: T.Tensor((T.ceildiv(N, group), T.ceildiv(K, group)), "float32")):
        with T.Kernel(T.ceildiv(N, bN), T.ceildiv(M, bM), threads=128) as (bx, by):
            a_s = T.alloc_shared((bM, bK), in_dtype)
            b_s = T.alloc_shared((bN, bK), in_dtype)
            c_s = T.alloc_shared((bM, bN), out_dtype)
            row_scale = T.alloc_shared((bM,), "float32")
            part = T.alloc_fragment((bM, bN), "float32")
            acc = T.alloc_fragment((bM, bN), "float32")
            T.clear(part)
            T.clear(acc)
            for k in T.Pipelined(T.ceildiv(K, bK), num_stages=4):
                T.copy(A[by * bM, k * bK], a_s)
                T.copy(B[bx * bN, k * bK], b_s)
                s_b = sb[bx * bN // group, k]
                for i in T.Parallel(bM):
                    if swap_scale_operands:
                        row_scale[i] = s_b * sa[by * bM + i, k]
                    else:
                        row_scale[i] = sa[by * bM + i, k] * s_b
                T.gemm(a_s, b_s, part, transpose_B=True)
                for i, j in T.Parallel(bM, bN):
                    acc[i, j] += part[i, j] * row_scale[i]
                T.clear(part)
            T.copy(acc, c_s)
            T.copy(c_s, C[by * bM, bx * bN])

    return prog


def splitk_atomic(M, N, K, bM=128, bN=128, bK=32, split_k=4, dtype="float16", out_dtype="float32", tile_form=False):
    """K cut over a third grid dimension, partial tiles added into C with T.atomic_add
    (the form of examples/gemm_splitk/example_tilelang_gemm_splitk.py:18-40 and its _vectorize_atomicadd variant)."""
    k_slice = K // split_k

    @T.prim_func
    def prog(A: T.Tensor((M, K), dtype), B: T.Tensor((K, N), dtype), C: T.Tensor((M, N), out_dtype)):
        with T.Kernel(T.ceildiv(N, bN), T.ceildiv(M, bM), split_k, threads=128) as (bx, by, bz):
            a_s = T.alloc_shared((bM, bK), dtype)
            b_s = T.alloc_shared((bK, bN), dtype)
            c_s = T.alloc_shared((bM, bN), out_dtype)
            acc = T.alloc_fragment((bM, bN), "float")
            T.clear(acc)
            for ko in T.Pipelined(T.ceildiv(k_slice, bK), num_stages=0):
                T.copy(A[by * bM, bz * k_slice + ko * bK], a_s)
                T.copy(B[bz * k_slice + ko * bK, bx * bN], b_s)
                T.gemm(a_s, b_s, acc)
            T.copy(acc, c_s)
            if tile_form:
                T.atomic_add(C[by * bM, bx * bN], c_s)
            else:
                for i, j in T.Parallel(bM, bN):
                    T.atomic_add(C[by * bM + i, bx * bN + j], c_s[i, j])

    return prog


def dynamic_mnk(bM=128, bN=128, bK=32, trans_A=False, trans_B=False, dtype="float16", out_dtype="float16", use_te_var=False):
    """All three extents are shape symbols (the form of examples/dynamic_shape/example_dynamic.py:19-52)."""
    if use_te_var:
        from tilelang import tvm
        M, N, K = tvm.te.var("m"), tvm.te.var("n"), tvm.te.var("k")
    else:
        M, N, K = T.dynamic("m"), T.dynamic("n"), T.dynamic("k")
    a_shape, b_shape = ((K, M) if trans_A else (M, K)), ((N, K) if trans_B else (K, N))
    a_tile, b_tile = ((bK, bM) if trans_A else (bM, bK)), ((bN, bK) if trans_B else (bK, bN))

    @T.prim_func
    def prog(A: T.Tensor(a_shape, dtype), B: T.Tensor(b_shape, dtype), C: T.Tensor((M, N), out_dtype)):
        with T.Kernel(T.ceildiv(N, bN), T.ceildiv(M, bM), threads=128) as (bx, by):
            a_s = T.alloc_shared(a_tile, dtype)
            b_s = T.alloc_shared(b_tile, dtype)
            acc = T.alloc_fragment((bM, bN), "float")
            T.clear(acc)
            for k in T.Pipelined(T.ceildiv(K, bK), num_stages=3):
                T.copy(A[(k * bK, by * bM) if trans_A else (by * bM, k * bK)], a_s)
                T.copy(B[(bx * bN, k * bK) if trans_B else (k * bK, bx * bN)], b_s)
                T.gemm(a_s, b_s, acc, trans_A, trans_B)
            T.copy(acc, C[by * bM, bx * bN])

    return prog


def warp_specialized(M, N, K, bM=128, bN=128, bK=64, dtype="float16"):
    """Hand-specialised producer / consumer warp groups with user barriers (the form of
    examples/warp_specialize/example_warp_specialize_gemm_copy_0_gemm_1.py:19-48)."""

    @T.prim_func
    def prog(A: T.Tensor[(M, K), dtype], B: T.Tensor[(K, N), dtype], C: T.Tensor[(M, N), dtype]):
        with T.Kernel(T.ceildiv(N, bN), T.ceildiv(M, bM), threads=256) as (bx, by):
            a_s = T.alloc_shared((bM, bK), dtype)
            b_s = T.alloc_shared((bK, bN), dtype)
            acc = T.alloc_fragment((bM, bN), "float")
            ready = T.alloc_barrier(arrive_count=128)
            done = T.alloc_barrier(arrive_count=128)
            with T.ws(1):
                T.clear(acc)
            for ko in T.Pipelined(T.ceildiv(K, bK), num_stages=0):
                with T.ws(0):
                    T.barrier_wait(done, (ko + 1) % 2)
                    T.copy(A[by * bM, ko * bK], a_s)
                    T.copy(B[ko * bK, bx * bN], b_s)
                    T.barrier_arrive(ready)
                with T.ws(1):
                    T.barrier_wait(ready, ko % 2)
                    T.gemm(a_s, b_s, acc)
                    T.barrier_arrive(done)
            with T.ws(1):
                T.copy(acc, C[by * bM, bx * bN])

    return prog


def fragment_operands(M, N, K, a_frag=True, b_frag=False, bM=128, bN=128, bK=32, trans_A=False, trans_B=True, dtype="float16"):
    """Operands handed to T.gemm as register fragments copied from their shared tiles: the gemm_rs / gemm_sr / gemm_rr
    forms (testing/python/tilelibrary/test_tilelang_tilelibrary_gemm.py:137-186 and the sr / rr siblings)."""
    a_shape, b_shape = ((K, M) if trans_A else (M, K)), ((N, K) if trans_B else (K, N))
    a_tile, b_tile = ((bK, bM) if trans_A else (bM, bK)), ((bN, bK) if trans_B else (bK, bN))

    @T.prim_func
    def prog(A: T.Tensor(a_shape, dtype), B: T.Tensor(b_shape, dtype), C: T.Tensor((M, N), dtype)):
        with T.Kernel(T.ceildiv(N, bN), T.ceildiv(M, bM), threads=128) as (bx, by):
            a_s = T.alloc_shared(a_tile, dtype)
            b_s = T.alloc_shared(b_tile, dtype)
            a_f = T.alloc_fragment(a_tile, dtype)
            b_f = T.alloc_fragment(b_tile, dtype)
            acc = T.alloc_fragment((bM, bN), "float")
            T.clear(acc)
            for k in T.Pipelined(T.ceildiv(K, bK), num_stages=2):
                T.copy(A[(k * bK, by * bM) if trans_A else (by * bM, k * bK)], a_s)
                T.copy(B[(bx * bN, k * bK) if trans_B else (k * bK, bx * bN)], b_s)
                if a_frag:
                    T.copy(a_s, a_f)
                if b_frag:
                    T.copy(b_s, b_f)
                T.gemm_v2(a_f if a_frag else a_s, b_f if b_frag else b_s, acc, trans_A, trans_B)
            T.copy(acc, C[by * bM, bx * bN])

    return prog
