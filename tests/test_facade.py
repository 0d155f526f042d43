"""CPU tests of the facade: tracing, GEMM-family recognition, JIT surface, error behaviour.
(`import tilelang` resolves to this repo's backend through compat/.)"""
import os
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "compat"))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

import tilelang  # noqa: E402
import tilelang.language as T  # noqa: E402
import tile_programs as P  # noqa: E402
from tilelang_mesh_b200.program import NotAGemmProgram  # noqa: E402

REF = "/root/reference"


def test_alias_is_this_backend():
    assert tilelang.__name__ == "tilelang_mesh_b200"
    assert T.gemm is T.gemm_v2  # default variant, tilelang/language/gemm.py:163
    import inspect
    sig = inspect.signature(T.gemm)
    assert list(sig.parameters) == ["A", "B", "C", "transpose_A", "transpose_B", "policy", "clear_accum", "k_pack",
                                    "wg_wait", "mbar"]
    assert [p.default for p in sig.parameters.values()][3:] == [False, False, T.GemmWarpPolicy.Square, False, 1, 0, None]
    assert list(inspect.signature(T.copy).parameters) == ["src", "dst", "coalesced_width", "disable_tma", "eviction_policy"]
    assert [int(x) for x in T.GemmWarpPolicy] == [0, 1, 2]


def test_recognize_benchmark_style_program():
    k = tilelang.compile(P.nt_fragment(8192, 8192, 8192, 128, 256, 64, "bfloat16"), out_idx=[2])
    p = k.program
    assert (p.M, p.N, p.K, p.in_dtype, p.out_dtype, p.trans_A, p.trans_B) == (8192, 8192, 8192, "bfloat16", "bfloat16", False, True)
    assert (p.a_param, p.b_param, p.c_param) == (0, 1, 2) and k.out_idx == [2] and not p.relu
    assert p.accumulator == "fragment" and p.rasterization["panel_size"] == 10
    plan = k.plan()
    assert plan["cta_group"] == 2 and plan["block_n"] == 512 and plan["umma_n"] == 256


def test_recognize_int8_and_float32_programs():
    p = tilelang.compile(P.nt_fragment(512, 1024, 768, 128, 128, 64, "int8", "int32", accum="int32"), out_idx=[2]).program
    assert (p.in_dtype, p.out_dtype, p.accum_dtype) == ("int8", "int32", "int32")
    p = tilelang.compile(P.nt_fragment(512, 1024, 768, 64, 128, 32, "float32"), out_idx=[2]).program
    assert (p.in_dtype, p.out_dtype, p.accum_dtype) == ("float32", "float32", "float32")
    with pytest.raises(NotImplementedError, match="int8 operands accumulate in int32"):
        tilelang.compile(P.nt_fragment(128, 128, 128, dtype="int8", out_dtype="int32"))


def test_recognize_quickstart_c1_descriptor():
    # configs[0] (C1): 1024^3 fp16 NN + relu, C passed by the caller (no out_idx)
    k = tilelang.compile(P.nn_relu(1024, 1024, 1024))
    p = k.program
    assert (p.M, p.N, p.K, p.trans_A, p.trans_B, p.relu, p.in_dtype) == (1024, 1024, 1024, False, False, True, "float16")
    assert k.out_idx == [] and [tuple(q.shape) for q in k.params] == [(1024, 1024)] * 3


def test_c1_descriptor_numerics_via_oracle():
    # plumbing check without a GPU: the recognised descriptor, evaluated by the ORACLE, reproduces
    # relu(A @ B) — i.e. the facade understood the program (the product path itself is CUDA-only)
    from oracle import gemm_ref as R
    p = tilelang.compile(P.nn_relu(64, 48, 32, 32, 16, 16)).program
    a, b = torch.randn(64, 32).half(), torch.randn(32, 48).half()
    got = R.c_oracle_gemm(a, b, torch.float16, p.trans_A, p.trans_B, p.alpha, p.relu)
    R.assert_allclose(got, torch.relu(a.float() @ b.float()).half(), 1e-2, 1e-2)


@pytest.mark.parametrize("ta,tb", [(False, True), (False, False), (True, True), (True, False)])
def test_recognize_tmem_program_all_layouts(ta, tb):
    k = tilelang.compile(P.tmem_accumulator(512, 768, 1024, trans_A=ta, trans_B=tb), out_idx=[2],
                         pass_configs={tilelang.PassConfigKey.TL_DISABLE_TMA_LOWER: True,
                                       tilelang.PassConfigKey.TL_DISABLE_WARP_SPECIALIZED: True})
    p = k.program
    assert (p.trans_A, p.trans_B, p.accumulator, p.num_stages) == (ta, tb, "tmem", 2)
    assert (p.M, p.N, p.K) == (512, 768, 1024)


def test_jit_factory_caches_and_out_idx_forms():
    calls = []

    @tilelang.jit(out_idx=-1)
    def factory(M, N, K, bM=128):
        calls.append(M)
        return P.nt_fragment(M, N, K, bM)

    k1, k2, k3 = factory(256, 256, 256), factory(256, 256, 256), factory(512, 256, 256)
    assert k1 is k2 and k1 is not k3 and calls == [256, 512]
    assert k1.out_idx == [2]
    with pytest.raises(ValueError, match="output index"):
        tilelang.compile(P.nt_fragment(128, 128, 128), out_idx=[3])
    with pytest.raises(ValueError, match="names an input"):
        tilelang.compile(P.nt_fragment(128, 128, 128), out_idx=[0])

    @tilelang.jit
    def bare(M):
        return P.nn_relu(M, M, M)

    assert bare(128).out_idx == []


def test_argument_validation_matches_reference_adapter():
    k = tilelang.compile(P.nt_fragment(128, 128, 128), out_idx=[2])
    a = torch.zeros(128, 128, dtype=torch.float16)
    with pytest.raises(ValueError, match="Expected 3 inputs"):
        k(a)
    with pytest.raises(ValueError, match="dtype mismatch"):
        k(a, a.bfloat16())
    with pytest.raises(ValueError, match="shape mismatch"):
        k(a, torch.zeros(64, 128, dtype=torch.float16))
    with pytest.raises(ValueError, match="CUDA device"):
        k(a, a)  # CPU tensors: no fallback


def test_outside_the_family_is_rejected_loudly():
    M = N = K = 256

    @T.prim_func
    def two_gemms(A: T.Tensor((M, K), "float16"), B: T.Tensor((N, K), "float16"), C: T.Tensor((M, N), "float16")):
        with T.Kernel(2, 2, threads=128) as (bx, by):
            a_s = T.alloc_shared((128, 32), "float16")
            b_s = T.alloc_shared((128, 32), "float16")
            acc = T.alloc_fragment((128, 128), "float")
            T.clear(acc)
            for k in T.Pipelined(8, num_stages=2):
                T.copy(A[by * 128, k * 32], a_s)
                T.copy(B[bx * 128, k * 32], b_s)
                T.gemm(a_s, b_s, acc, transpose_B=True)
                T.gemm(a_s, b_s, acc, transpose_B=True)
            T.copy(acc, C[by * 128, bx * 128])

    with pytest.raises(NotImplementedError, match="exactly one T.gemm"):
        tilelang.compile(two_gemms)

    @T.prim_func
    def never_cleared(A: T.Tensor((M, K), "float16"), B: T.Tensor((N, K), "float16"), C: T.Tensor((M, N), "float16")):
        with T.Kernel(2, 2, threads=128) as (bx, by):
            a_s = T.alloc_shared((128, 32), "float16")
            b_s = T.alloc_shared((128, 32), "float16")
            acc = T.alloc_fragment((128, 128), "float")
            for k in T.Pipelined(8, num_stages=2):
                T.copy(A[by * 128, k * 32], a_s)
                T.copy(B[bx * 128, k * 32], b_s)
                T.gemm(a_s, b_s, acc, transpose_B=True)
            T.copy(acc, C[by * 128, bx * 128])

    with pytest.raises(NotAGemmProgram, match="never cleared"):
        tilelang.compile(never_cleared)

    @T.prim_func
    def wrong_k(A: T.Tensor((M, K), "float16"), B: T.Tensor((N, K), "float16"), C: T.Tensor((M, N), "float16")):
        with T.Kernel(2, 2, threads=128) as (bx, by):
            a_s = T.alloc_shared((128, 32), "float16")
            b_s = T.alloc_shared((128, 64), "float16")
            acc = T.alloc_fragment((128, 128), "float")
            T.clear(acc)
            for k in T.Pipelined(8, num_stages=2):
                T.copy(A[by * 128, k * 32], a_s)
                T.gemm(a_s, b_s, acc, transpose_B=True)
            T.copy(acc, C[by * 128, bx * 128])

    with pytest.raises(AssertionError, match="K shape check failed"):  # language/gemm.py:78-81
        tilelang.compile(wrong_k)

    with pytest.raises(NotImplementedError, match="sm_100a code only"):
        tilelang.compile(P.nt_fragment(128, 128, 128), target="hip")

    @T.prim_func
    def cpu_kernel(A: T.Tensor((M, K), "float16"), B: T.Tensor((K, N), "float16"), C: T.Tensor((M, N), "float16")):
        with T.Kernel(2, 2, is_cpu=True) as (bx, by):
            pass

    with pytest.raises(NotImplementedError, match="no CPU path"):
        tilelang.compile(cpu_kernel)


def test_scaled_epilogue_and_partial_k_loop_rejected():
    M = N = K = 256

    @T.prim_func
    def scaled(A: T.Tensor((M, K), "float8_e4m3"), B: T.Tensor((N, K), "float8_e4m3"), C: T.Tensor((M, N), "bfloat16")):
        with T.Kernel(2, 2, threads=128) as (bx, by):
            a_s = T.alloc_shared((128, 64), "float8_e4m3")
            b_s = T.alloc_shared((128, 64), "float8_e4m3")
            acc = T.alloc_fragment((128, 128), "float")
            T.clear(acc)
            for k in T.Pipelined(4, num_stages=2):
                T.copy(A[by * 128, k * 64], a_s)
                T.copy(B[bx * 128, k * 64], b_s)
                T.gemm(a_s, b_s, acc, transpose_B=True)
            for i, j in T.Parallel(128, 128):
                acc[i, j] = acc[i, j] * 0.25
            T.copy(acc, C[by * 128, bx * 128])

    p = tilelang.compile(scaled, out_idx=[2]).program
    assert p.alpha == 0.25 and p.in_dtype == "float8_e4m3" and p.out_dtype == "bfloat16"

    @T.prim_func
    def half_k(A: T.Tensor((M, K), "float16"), B: T.Tensor((N, K), "float16"), C: T.Tensor((M, N), "float16")):
        with T.Kernel(2, 2, threads=128) as (bx, by):
            a_s = T.alloc_shared((128, 32), "float16")
            b_s = T.alloc_shared((128, 32), "float16")
            acc = T.alloc_fragment((128, 128), "float")
            T.clear(acc)
            for k in T.Pipelined(4, num_stages=2):
                T.copy(A[by * 128, k * 32], a_s)
                T.copy(B[bx * 128, k * 32], b_s)
                T.gemm(a_s, b_s, acc, transpose_B=True)
            T.copy(acc, C[by * 128, bx * 128])

    with pytest.raises(NotAGemmProgram, match="K loop runs 4"):
        tilelang.compile(half_k)


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference tree not mounted (only in the build container)")
@pytest.mark.parametrize("rel,cut,call", [
    ("examples/quickstart.py", "M = 1024", "matmul(1024, 1024, 1024, 128, 128, 32)"),
    ("examples/gemm/example_gemm.py", "def main", "matmul(1024, 1024, 1024, 128, 128, 32)"),
    ("examples/gemm_sm100/gemm_tcgen5mma.py", "print(jit_kernel.get_kernel_source())", "jit_kernel"),
    ("examples/gemm/example_gemm_schedule.py", "def main", "matmul(1024, 1024, 1024, 128, 128, 32)"),
    ("examples/gemm/example_gemm_autotune.py", "\ndef main(M", "matmul(1024, 1024, 1024, 128, 256, 32, 0, 128, True)"),
    ("examples/gemm/example_gemm_autotune.py", "\ndef main(M", "get_best_config(1024, 1024, 1024, False).kernel"),
    ("examples/gemm/example_gemm_persistent.py", "def ref_program",
     "matmul_non_persistent(1024, 1024, 1024, 128, 256, 32, 128, 3)"),
    ("examples/deepseek_deepgemm/example_deepgemm_fp8_2xAcc.py", "def ceildiv",
     'tl_gemm(1024, 1024, 8192, 64, "float8_e4m3", "bfloat16", "float32")'),
    ("examples/gemm_splitk/example_tilelang_gemm_splitk.py", "def main()", "matmul(1024, 1024, 1024, 128, 128, 32, 4)"),
    ("examples/gemm_splitk/example_tilelang_gemm_splitk_vectorize_atomicadd.py", "def main()",
     "matmul(1024, 1024, 1024, 128, 128, 32, 4)"),
    ("examples/warp_specialize/example_warp_specialize_gemm_copy_0_gemm_1.py", "\ndef main(",
     "matmul_warp_specialize_copy_0_gemm_1(1024, 1024, 1024, 128, 128, 64)"),
    ("examples/warp_specialize/example_warp_specialize_gemm_copy_1_gemm_0.py", "\ndef main(",
     "matmul_warp_specialize_copy_1_gemm_0(1024, 1024, 1024, 128, 128, 64)"),
    ("examples/warp_specialize/example_warp_specialize_gemm_softpipe_stage2.py", "\ndef main(",
     "matmul(1024, 1024, 1024, 128, 128, 64)"),
    ("examples/dynamic_shape/example_dynamic.py", "\ndef matmul_dynamic(",
     'matmul_dynamic_mnk(128, 128, 32, False, False, "float16", "float16", "float32", 3, 128)'),
    ("testing/python/tilelibrary/test_tilelang_tilelibrary_gemm.py", "\nif __name__", 'tilelang.compile(matmul_rs(512, 1024, 768, '
     '128, 256, 32, False, True, "float16", "float16", "float32", 2, 128), out_idx=[2])'),
    ("testing/python/tilelibrary/test_tilelang_tilelibrary_gemm.py", "\nif __name__", 'tilelang.compile(matmul_rr(512, 1024, 768, '
     '128, 256, 32, True, False, "bfloat16", "bfloat16", "float32", 2, 128), out_idx=[2])'),
    ("testing/python/kernel/test_tilelang_kernel_gemm.py", "\nif __name__", 'tilelang.compile(matmul(512, 1024, 768, 128, 128, 32, '
     'False, True, "float32", "float32", "float32", 2, 128), out_idx=[2])'),
    ("testing/python/kernel/test_tilelang_kernel_fp8_gemm.py", "\nif __name__", 'tilelang.compile(matmul_nt(1024, 1024, 1024, 128, '
     '128, 64, "float8_e4m3", "bfloat16", "float32"), out_idx=[-1])'),
    ("benchmark/matmul/benchmark_matmul.py", 'if __name__ == "__main__":', "matmul(8192, 8192, 8192, False)"),
    ("benchmark/matmul_fp8/benchmark_matmul.py", 'if __name__ == "__main__":', "matmul(8192, 8192, 8192, False)"),
])
def test_reference_programs_trace_unmodified(rel, cut, call):
    # Drop-in check in the build container: the reference's own program text, executed verbatim
    # against this facade, yields a GEMM descriptor (nothing is copied into the repo).
    src = open(os.path.join(REF, rel)).read().split(cut)[0]
    ns = {"__name__": "ref_program"}
    exec(compile(src, rel, "exec"), ns)
    kern = eval(call, ns)
    p = kern.program
    assert p.M > 0 and p.accum_dtype == "float32"
    assert kern.plan()["grid"] >= 1
    assert "tcgen05" in kern.get_kernel_source()


def test_elementwise_staging_loop_is_a_copy():
    k = tilelang.compile(P.elementwise_staging(512, 384, 256), out_idx=[2])
    p = k.program
    assert (p.M, p.N, p.K, p.trans_A, p.trans_B) == (512, 384, 256, False, False)

    M = N = K = 256

    @T.prim_func
    def transposing_fill(A: T.Tensor((M, K), "float16"), B: T.Tensor((K, N), "float16"), C: T.Tensor((M, N), "float16")):
        with T.Kernel(2, 2, threads=128) as (bx, by):
            a_s = T.alloc_shared((128, 32), "float16")
            b_s = T.alloc_shared((32, 128), "float16")
            acc = T.alloc_fragment((128, 128), "float")
            T.clear(acc)
            for k in T.Pipelined(8, num_stages=2):
                T.copy(A[by * 128, k * 32], a_s)
                for kk, j in T.Parallel(32, 128):
                    b_s[kk, j] = B[k * 32 + j, bx * 128 + kk]
                T.gemm(a_s, b_s, acc)
            T.copy(acc, C[by * 128, bx * 128])

    with pytest.raises(NotAGemmProgram, match="must be origin"):
        tilelang.compile(transposing_fill)


def test_persistent_primitive_program():
    k = tilelang.compile(P.persistent_primitive(1024, 2048, 512), out_idx=[-1])
    p = k.program
    assert (p.M, p.N, p.K, p.block_M, p.block_N) == (1024, 2048, 512, 128, 256)
    # tile domain must cover the output
    with pytest.raises(NotAGemmProgram, match="grid dimension"):
        @T.prim_func
        def short(A: T.Tensor((1024, 64), "float16"), B: T.Tensor((64, 1024), "float16"), C: T.Tensor((1024, 1024), "float16")):
            with T.Kernel(8, threads=128) as (block_id):
                a_s = T.alloc_shared((128, 64), "float16")
                b_s = T.alloc_shared((64, 128), "float16")
                acc = T.alloc_fragment((128, 128), "float")
                for tx, ty in T.Persistent([4, 8], 8, block_id):
                    T.clear(acc)
                    for k in T.Pipelined(1, num_stages=1):
                        T.copy(A[tx * 128, k * 64], a_s)
                        T.copy(B[k * 64, ty * 128], b_s)
                        T.gemm(a_s, b_s, acc)
                    T.copy(acc, C[tx * 128, ty * 128])
        tilelang.compile(short)


def test_fp16_accumulate_programs_are_widened():
    # examples/gemm_fp8/example_tilelang_gemm_fp8_sm100.py:76 runs accum_dtype float16 for fp8 operands
    k = tilelang.compile(P.tmem_accumulator(256, 256, 256, 64, 256, 32, in_dtype="float8_e4m3", out_dtype="float16",
                                            accum_dtype="float16"), out_idx=[2])
    assert k.program.accum_dtype == "float16" and any("widened" in n for n in k.program.notes)
    with pytest.raises(NotAGemmProgram, match="accumulator dtype"):
        tilelang.compile(P.tmem_accumulator(256, 256, 256, in_dtype="bfloat16", out_dtype="float16", accum_dtype="float16"))


def test_autotuner_builder_form_and_carver_names():
    from tilelang.autotuner import AutoTuner
    from tilelang.carver.arch import CUDA
    from tilelang.carver.roller.rasterization import NoRasterization  # noqa: F401
    from tilelang.carver.template import MatmulTemplate

    def kernel(bM=None, bN=None, bK=None):
        return P.nt_fragment(512, 512, 256, bM, bN, bK)

    cfgs = [dict(bM=128, bN=128, bK=32), dict(bM=128, bN=256, bK=64)]
    tuner = AutoTuner.from_kernel(kernel=kernel, configs=cfgs).set_compile_args(out_idx=[-1], target="auto")
    tuner = tuner.set_profile_args(ref_prog=lambda a, b: a @ b.T, skip_check=True)
    if not torch.cuda.is_available():
        res = tuner.run(warmup=1, rep=2)
        assert res.config["bM"] == 128 and res.kernel.program.M == 512 and res.latency is None
    with pytest.raises(NotImplementedError, match="plan"):
        MatmulTemplate(M=512, N=512, K=256, in_dtype="float16", out_dtype="float16", accum_dtype="float")
    with pytest.raises(NotImplementedError):
        CUDA("cuda")


def test_blockscaled_2xacc_program_family():
    from tilelang_mesh_b200.program import BlockScaledProgram
    for bn, swap in ((128, False), (64, True), (16, False)):
        k = tilelang.compile(P.blockscaled_2xacc(512, 384, 1024, bn, "float32", swap_scale_operands=swap))
        p = k.program
        assert isinstance(p, BlockScaledProgram) and (p.M, p.N, p.K) == (512, 384, 1024)
        assert (p.a_param, p.b_param, p.c_param, p.sa_param, p.sb_param) == (0, 1, 2, 3, 4)
        assert [tuple(kp.shape) for kp in k.params] == [(512, 1024), (384, 1024), (512, 384), (512, 8), (3, 8)]
        assert "tcgen05" in k.get_kernel_source() and k.plan()["block_k"] == 128
    # the scale group is the K block: any other block_K changes which products share a scale
    with pytest.raises(NotAGemmProgram, match="block_K must equal the scale group"):
        tilelang.compile(P.blockscaled_2xacc(512, 384, 1024, 128, bK=64))
    with pytest.raises(NotAGemmProgram, match="block_N must divide"):
        tilelang.compile(P.blockscaled_2xacc(512, 512, 1024, 256))
    with pytest.raises(NotAGemmProgram, match="fp8"):
        tilelang.compile(P.blockscaled_2xacc(512, 384, 1024, 128, in_dtype="float16"))
    with pytest.raises(NotAGemmProgram, match="scales_b must be indexed"):
        tilelang.compile(P.blockscaled_2xacc(512, 384, 1024, 128, group=64))


def test_splitk_atomic_add_programs():
    for tile_form in (False, True):
        k = tilelang.compile(P.splitk_atomic(512, 384, 2048, split_k=4, tile_form=tile_form))
        p = k.program
        assert (p.M, p.N, p.K, p.accumulate, p.split_k, p.out_dtype) == (512, 384, 2048, True, 4, "float32")
    with pytest.raises(NotAGemmProgram, match="tile K exactly"):
        tilelang.compile(P.splitk_atomic(512, 384, 2000, split_k=3))  # 2000 // 3 slices drop the tail of K

    M = N = 256

    @T.prim_func
    def sliced_but_overwriting(A: T.Tensor((M, 512), "float16"), B: T.Tensor((512, N), "float16"), C: T.Tensor((M, N), "float32")):
        with T.Kernel(2, 2, 2, threads=128) as (bx, by, bz):
            a_s = T.alloc_shared((128, 32), "float16")
            b_s = T.alloc_shared((32, 128), "float16")
            acc = T.alloc_fragment((128, 128), "float")
            T.clear(acc)
            for ko in T.Pipelined(8, num_stages=2):
                T.copy(A[by * 128, bz * 256 + ko * 32], a_s)
                T.copy(B[bz * 256 + ko * 32, bx * 128], b_s)
                T.gemm(a_s, b_s, acc)
            T.copy(acc, C[by * 128, bx * 128])

    with pytest.raises(NotAGemmProgram, match="must add its partial tiles"):
        tilelang.compile(sliced_but_overwriting)


def test_dynamic_shape_and_warp_specialized_programs():
    for te_var in (False, True):
        k = tilelang.compile(P.dynamic_mnk(use_te_var=te_var, trans_B=True))
        assert k._dynamic and [v.name for v in k._symbols] == ["m", "k", "n"]
        assert k.program.trans_B and not k.program.trans_A and k.program.block_K == 32
        assert all(kp.is_dynamic for kp in k.params)
    k = tilelang.compile(P.warp_specialized(512, 768, 1024), out_idx=[2])
    assert (k.program.M, k.program.N, k.program.K) == (512, 768, 1024)


@pytest.mark.parametrize("a_frag,b_frag", [(True, False), (False, True), (True, True)], ids=["rs", "sr", "rr"])
def test_fragment_operand_forms(a_frag, b_frag):
    k = tilelang.compile(P.fragment_operands(512, 384, 256, a_frag, b_frag), out_idx=[2])
    assert (k.program.M, k.program.N, k.program.K) == (512, 384, 256)
    assert sum("fragment copy" in n for n in k.program.notes) == int(a_frag) + int(b_frag)


def test_fragment_operand_must_be_a_plain_copy_of_a_shared_tile():
    @T.prim_func
    def stale(A: T.Tensor((256, 256), "float16"), B: T.Tensor((256, 256), "float16"), C: T.Tensor((256, 256), "float16")):
        with T.Kernel(2, 2, threads=128) as (bx, by):
            a_s = T.alloc_shared((128, 32), "float16")
            b_s = T.alloc_shared((128, 32), "float16")
            a_f = T.alloc_fragment((128, 32), "float16")
            acc = T.alloc_fragment((128, 128), "float")
            T.clear(acc)
            T.copy(a_s, a_f)  # outside the K loop: every iteration would multiply the same stale tile
            for k in T.Pipelined(8, num_stages=2):
                T.copy(A[by * 128, k * 32], a_s)
                T.copy(B[bx * 128, k * 32], b_s)
                T.gemm(a_f, b_s, acc, transpose_B=True)
            T.copy(acc, C[by * 128, bx * 128])

    with pytest.raises(NotAGemmProgram, match="not a plain copy of a shared tile"):
        tilelang.compile(stale)


def test_writes_that_change_the_meaning_are_rejected():
    """Programs that look like the family but compute something else must not be folded into a plain GEMM."""
    M = N = K = 256

    @T.prim_func
    def scale_inside_k_loop(A: T.Tensor((M, K), "float16"), B: T.Tensor((N, K), "float16"), C: T.Tensor((M, N), "float16")):
        with T.Kernel(2, 2, threads=128) as (bx, by):
            a_s = T.alloc_shared((128, 32), "float16")
            b_s = T.alloc_shared((128, 32), "float16")
            acc = T.alloc_fragment((128, 128), "float")
            T.clear(acc)
            for k in T.Pipelined(8, num_stages=2):
                T.copy(A[by * 128, k * 32], a_s)
                T.copy(B[bx * 128, k * 32], b_s)
                T.gemm(a_s, b_s, acc, transpose_B=True)
                for i, j in T.Parallel(128, 128):  # once per K block: a geometric rescaling, not alpha = 0.5
                    acc[i, j] = acc[i, j] * 0.5
            T.copy(acc, C[by * 128, bx * 128])

    with pytest.raises(NotAGemmProgram, match="inside the K loop"):
        tilelang.compile(scale_inside_k_loop)

    @T.prim_func
    def operand_tile_scaled_in_place(A: T.Tensor((M, K), "float16"), B: T.Tensor((N, K), "float16"), C: T.Tensor((M, N), "float16")):
        with T.Kernel(2, 2, threads=128) as (bx, by):
            a_s = T.alloc_shared((128, 32), "float16")
            b_s = T.alloc_shared((128, 32), "float16")
            acc = T.alloc_fragment((128, 128), "float")
            T.clear(acc)
            for k in T.Pipelined(8, num_stages=2):
                T.copy(A[by * 128, k * 32], a_s)
                T.copy(B[bx * 128, k * 32], b_s)
                for i, j in T.Parallel(128, 32):  # result is 2*A@B, not A@B
                    a_s[i, j] = a_s[i, j] * 2
                T.gemm(a_s, b_s, acc, transpose_B=True)
            T.copy(acc, C[by * 128, bx * 128])

    with pytest.raises(NotAGemmProgram, match="modified element-wise"):
        tilelang.compile(operand_tile_scaled_in_place)

    @T.prim_func
    def work_after_the_store(A: T.Tensor((M, K), "float16"), B: T.Tensor((N, K), "float16"), C: T.Tensor((M, N), "float16")):
        with T.Kernel(2, 2, threads=128) as (bx, by):
            a_s = T.alloc_shared((128, 32), "float16")
            b_s = T.alloc_shared((128, 32), "float16")
            acc = T.alloc_fragment((128, 128), "float")
            T.clear(acc)
            for k in T.Pipelined(8, num_stages=2):
                T.copy(A[by * 128, k * 32], a_s)
                T.copy(B[bx * 128, k * 32], b_s)
                T.gemm(a_s, b_s, acc, transpose_B=True)
            T.copy(acc, C[by * 128, bx * 128])
            T.clear(acc)
            T.copy(acc, C[by * 128, bx * 128])  # overwrites the product with zeros

    with pytest.raises(NotAGemmProgram, match="unexpected"):
        tilelang.compile(work_after_the_store)

    @T.prim_func
    def fill_inside_k_loop(A: T.Tensor((M, K), "float16"), B: T.Tensor((N, K), "float16"), C: T.Tensor((M, N), "float16")):
        with T.Kernel(2, 2, threads=128) as (bx, by):
            a_s = T.alloc_shared((128, 32), "float16")
            b_s = T.alloc_shared((128, 32), "float16")
            acc = T.alloc_fragment((128, 128), "float")
            T.clear(acc)
            for k in T.Pipelined(8, num_stages=2):
                T.copy(A[by * 128, k * 32], a_s)
                T.copy(B[bx * 128, k * 32], b_s)
                T.fill(b_s, 1.0)
                T.gemm(a_s, b_s, acc, transpose_B=True)
            T.copy(acc, C[by * 128, bx * 128])

    with pytest.raises(NotAGemmProgram, match="unexpected ClearOp inside the K loop"):
        tilelang.compile(fill_inside_k_loop)
