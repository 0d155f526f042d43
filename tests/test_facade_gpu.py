"""GPU tests of the drop-in surface: tile programs compiled with `tilelang.jit` / `tilelang.compile`
(this repo's facade) run on the tcgen05 path and match the fp32 oracle the way the reference's own
tests demand (Profiler.assert_allclose atol=rtol=1e-2, <=1 % mismatches)."""
import os
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "compat"))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

import tilelang  # noqa: E402
import tile_programs as P  # noqa: E402
from oracle import gemm_ref as R  # noqa: E402
from tilelang_mesh_b200 import _abi  # noqa: E402

pytestmark = pytest.mark.gpu


def test_quickstart_style_c1_on_gpu():
    # C1 program (1024^3 fp16 NN + relu) through the CUDA path; caller passes C (no out_idx)
    kern = tilelang.compile(P.nn_relu(1024, 1024, 1024))
    a = torch.randn(1024, 1024, device="cuda", dtype=torch.float16)
    b = torch.randn(1024, 1024, device="cuda", dtype=torch.float16)
    c = torch.empty(1024, 1024, device="cuda", dtype=torch.float16)
    assert kern(a, b, c) is None
    torch.testing.assert_close(c, torch.relu(a @ b), rtol=1e-2, atol=1e-2)  # examples/quickstart.py:71-74
    prof = kern.get_profiler(tensor_supply_type=tilelang.TensorSupplyType.Normal)
    assert prof.do_bench() > 0


def test_sm100_example_style_program():
    # examples/gemm_sm100/gemm_tcgen5mma.py: 4096x4096x8192 bf16, tmem accumulator, out_idx=[2]
    M, N, K = 4096, 4096, 8192
    kern = tilelang.compile(P.tmem_accumulator(M, N, K), out_idx=[2], target="cuda",
                            pass_configs={tilelang.PassConfigKey.TL_DISABLE_TMA_LOWER: True,
                                          tilelang.PassConfigKey.TL_DISABLE_WARP_SPECIALIZED: True})
    a = torch.randn(M, K, device="cuda", dtype=torch.bfloat16)
    b = torch.randn(N, K, device="cuda", dtype=torch.bfloat16)
    c = kern(a, b)
    ref = (a.to(torch.float) @ b.T.to(torch.float)).to(torch.bfloat16)
    torch.testing.assert_close(c, ref, rtol=1e-2, atol=1e-2)
    lat = kern.get_profiler().do_bench()
    assert 2 * M * N * K / (lat / 1e3) / 1e12 > 300  # far above the mma.sync class of kernels


@pytest.mark.parametrize("ta,tb", [(False, False), (True, True), (True, False)], ids=["nn", "tn", "tt"])
def test_tmem_program_transposes(ta, tb):
    kern = tilelang.compile(P.tmem_accumulator(512, 768, 1024, trans_A=ta, trans_B=tb), out_idx=[2])
    prof = kern.get_profiler(tilelang.TensorSupplyType.Normal)

    def ref_program(A, B):  # maint/gemm_v2/correctness_evaluation_tcgen05.py:84-95
        A = A.T if ta else A
        B = B.T if tb else B
        return torch.matmul(A.float(), B.float()).to(torch.bfloat16)

    prof.assert_allclose(ref_program, atol=1e-2, rtol=1e-2)
    prof.assert_consistent(repeat=5)


def test_benchmark_style_program_with_autotune():
    from tilelang.autotuner import autotune

    def configs(args, kwargs):
        return [dict(bM=bm, bN=bn, bK=32, stages=s) for bm in (64, 128) for bn in (128, 256) for s in (2, 3)]

    @autotune(configs=configs, warmup=3, rep=20, ref_prog=lambda A, B: A @ B.T)
    @tilelang.jit(out_idx=[2])
    def matmul(M, N, K, bM=None, bN=None, bK=None, stages=None):
        return P.nt_fragment(M, N, K, bM, bN, bK, "float16", stages=stages)

    best = matmul(4096, 4096, 4096)
    assert best.latency > 0 and best.ref_latency > 0 and "backend_plan" in best.config
    tflops = 2 * 4096**3 / best.latency * 1e-9
    assert tflops > 300
    a = torch.randn(4096, 4096, device="cuda", dtype=torch.float16)
    b = torch.randn(4096, 4096, device="cuda", dtype=torch.float16)
    R.assert_allclose(best(a, b), R.matmul_fp32(a, b).half(), 1e-2, 1e-2)


def test_fp8_program_calc_diff():
    # testing/python/kernel/test_tilelang_kernel_fp8_gemm.py:38-60: 1024^3 e4m3, calc_diff < 1e-3
    kern = tilelang.compile(P.nt_fragment(1024, 1024, 1024, 128, 128, 64, "float8_e4m3", "float32"), out_idx=[2])
    a = torch.randn(1024, 1024, device="cuda").to(torch.float8_e4m3fn)
    b = torch.randn(1024, 1024, device="cuda").to(torch.float8_e4m3fn)
    c = kern(a, b)
    ref = a.to(torch.float32) @ b.to(torch.float32).T
    assert R.calc_diff(c, ref) < 1e-3


def test_reference_dtype_cases_f32_and_int8():
    # testing/python/kernel/test_tilelang_kernel_gemm.py: test_gemm_f32f32f32_nt / test_gemm_i8i8i32_nt at 512x1024x768
    kf = tilelang.compile(P.nt_fragment(512, 1024, 768, 64, 128, 32, "float32"), out_idx=[2])
    a = torch.randn(512, 768, device="cuda")
    b = torch.randn(1024, 768, device="cuda")
    R.assert_allclose(kf(a, b), a @ b.T, 1e-2, 1e-2)
    ki = tilelang.compile(P.nt_fragment(512, 1024, 768, 128, 128, 64, "int8", "int32", accum="int32"), out_idx=[2])
    a8 = torch.randint(-128, 128, (512, 768), device="cuda", dtype=torch.int8)
    b8 = torch.randint(-128, 128, (1024, 768), device="cuda", dtype=torch.int8)
    assert torch.equal(ki(a8, b8).cpu(), a8.cpu().int() @ b8.cpu().int().T)


def test_schedule_and_persistent_example_forms():
    # examples/gemm/example_gemm_schedule.py (element-wise B staging) and example_gemm_persistent.py (T.Persistent)
    for prog, dt in ((P.elementwise_staging(1024, 768, 512), torch.float16),
                     (P.persistent_primitive(1024, 2048, 512), torch.bfloat16)):
        kern = tilelang.compile(prog, out_idx=[-1])
        p = kern.program
        a = torch.randn(p.M, p.K, device="cuda", dtype=dt)
        b = torch.randn(p.K, p.N, device="cuda", dtype=dt)
        torch.testing.assert_close(kern(a, b), a @ b, rtol=1e-2, atol=1e-2)


def test_fp8_fp16_accumulate_program_widened():
    # examples/gemm_fp8/example_tilelang_gemm_fp8_sm100.py:75-118: fp8 operands, accum/out float16
    M, N, K = 1024, 1024, 2048
    for fp8, tdt in (("float8_e4m3", torch.float8_e4m3fn), ("float8_e5m2", torch.float8_e5m2)):
        kern = tilelang.compile(P.tmem_accumulator(M, N, K, 64, 256, 32, in_dtype=fp8, out_dtype="float16",
                                                   accum_dtype="float16", stages=2), out_idx=[2])
        a = torch.randn(M, K, device="cuda", dtype=torch.float16).to(tdt)
        b = torch.randn(N, K, device="cuda", dtype=torch.float16).to(tdt)
        c = kern(a, b)
        ref = (a.to(torch.half) @ b.T.to(torch.half)).float()
        assert c.dtype == torch.float16 and float(R.calc_diff(c.float(), ref)) < 1e-3


def test_autotuner_builder_form():
    from tilelang.autotuner import AutoTuner

    def kernel(bM=None, bN=None, bK=None):
        return P.nt_fragment(1024, 1024, 512, bM, bN, bK)

    res = AutoTuner.from_kernel(kernel=kernel, configs=[dict(bM=128, bN=128, bK=32)]).set_compile_args(
        out_idx=[-1], target="auto").set_profile_args(ref_prog=lambda a, b: a @ b.T, skip_check=False).run(warmup=1, rep=3)
    assert res.latency > 0 and res.ref_latency > 0 and res.config["bM"] == 128 and "backend_plan" in res.config
    a = torch.randn(1024, 512, device="cuda", dtype=torch.float16)
    b = torch.randn(1024, 512, device="cuda", dtype=torch.float16)
    torch.testing.assert_close(res.kernel(a, b), a @ b.T, rtol=1e-2, atol=1e-2)


def test_blockscaled_2xacc_program_through_the_facade():
    # examples/deepseek_deepgemm/example_deepgemm_fp8_2xAcc.py:139-172: kernel(A_fp8, B_fp8, C, A_scale, B_scale)
    M, N, K = 1024, 1024, 8192
    for bn, out, tdt in ((128, "bfloat16", torch.bfloat16), (32, "float32", torch.float32)):
        kern = tilelang.compile(P.blockscaled_2xacc(M, N, K, bn, out))
        torch.manual_seed(42)
        a = torch.randn(M, K, device="cuda").to(torch.float8_e4m3fn)
        b = torch.randn(N, K, device="cuda").to(torch.float8_e4m3fn)
        sa = torch.rand(M, K // 128, device="cuda") + 0.5
        sb = torch.rand(N // 128, K // 128, device="cuda") + 0.5
        c = torch.zeros(M, N, device="cuda", dtype=tdt)
        assert kern(a, b, c, sa, sb) is None
        assert R.calc_diff(c.float(), R.blockscaled_fp32(a, b, sa, sb)) < 1e-5  # reference bar: < 1e-3
        assert kern.get_profiler().do_bench(warmup=5, rep=10) > 0


def test_splitk_atomic_add_program_accumulates_into_c():
    # examples/gemm_splitk: c = zeros; kernel(a, b, c); c == a @ b.  A second call adds the product once more.
    M, N, K = 1024, 768, 4096
    for tile_form in (False, True):
        kern = tilelang.compile(P.splitk_atomic(M, N, K, split_k=4, tile_form=tile_form))
        a = torch.randint(-3, 4, (M, K), device="cuda").half()
        b = torch.randint(-3, 4, (K, N), device="cuda").half()
        c = torch.zeros(M, N, device="cuda", dtype=torch.float32)
        assert kern(a, b, c) is None
        want = a.float() @ b.float()
        assert torch.equal(c, want)
        kern(a, b, c)
        assert torch.equal(c, 2 * want)


def test_dynamic_shapes_bound_per_call():
    # examples/dynamic_shape/example_dynamic.py: one kernel object, any (m, n, k); a plan is resolved per shape
    for ta, tb in ((False, False), (False, True), (True, True)):
        kern = tilelang.compile(P.dynamic_mnk(trans_A=ta, trans_B=tb))
        for m, n, k in ((1024, 1024, 1024), (300, 520, 136), (4096, 256, 2048)):
            a = torch.randint(-3, 4, (k, m) if ta else (m, k), device="cuda").half()
            b = torch.randint(-3, 4, (n, k) if tb else (k, n), device="cuda").half()
            c = torch.zeros(m, n, device="cuda", dtype=torch.float16)
            kern(a, b, c)
            want = ((a.t() if ta else a).float() @ (b.t() if tb else b).float()).half()
            assert torch.equal(c, want)
        assert len(kern._programs) == 3
        with pytest.raises(ValueError, match="shape mismatch"):  # k of A and B disagree
            kern(torch.zeros((64, 128) if ta else (128, 64), device="cuda").half(),
                 torch.zeros((128, 96) if tb else (96, 128), device="cuda").half(), torch.zeros(128, 128, device="cuda").half())
    kern = tilelang.compile(P.dynamic_mnk(trans_B=True), out_idx=[2])  # output allocated from the bound symbols
    a = torch.randn(200, 264, device="cuda").half()
    b = torch.randn(96, 264, device="cuda").half()
    torch.testing.assert_close(kern(a, b), a @ b.t(), rtol=1e-2, atol=1e-2)


def test_warp_specialized_program():
    kern = tilelang.compile(P.warp_specialized(1024, 768, 512), out_idx=[2])
    a = torch.randn(1024, 512, device="cuda").half()
    b = torch.randn(512, 768, device="cuda").half()
    torch.testing.assert_close(kern(a, b), a @ b, rtol=1e-2, atol=1e-2)  # example's own check (:67)


@pytest.mark.parametrize("a_frag,b_frag", [(True, False), (False, True), (True, True)], ids=["rs", "sr", "rr"])
@pytest.mark.parametrize("ta,tb", [(False, True), (True, False)], ids=["nt", "tn"])
def test_fragment_operand_forms_on_gpu(a_frag, b_frag, ta, tb):
    # the reference's run_gemm_rs / _sr / _rr cases (test_tilelang_tilelibrary_gemm.py), checked the way they are:
    # profiler.assert_allclose against torch.matmul of the (transposed) inputs, rtol = atol = 1e-2
    kern = tilelang.compile(P.fragment_operands(512, 1024, 768, a_frag, b_frag, 128, 256, 32, ta, tb), out_idx=[2])

    def ref_program(A, B):
        A = A.T if ta else A
        B = B.T if tb else B
        return torch.matmul(A.float(), B.float()).half()

    kern.get_profiler(tilelang.TensorSupplyType.Normal).assert_allclose(ref_program, atol=1e-2, rtol=1e-2)


def test_kernel_source_and_multi_stream():
    kern = tilelang.compile(P.nt_fragment(512, 512, 512, dtype="bfloat16"), out_idx=[2])
    assert "tcgen05.mma" in kern.get_kernel_source(kernel_only=False)
    a = torch.randn(512, 512, device="cuda", dtype=torch.bfloat16)
    b = torch.randn(512, 512, device="cuda", dtype=torch.bfloat16)
    outs = []
    streams = [torch.cuda.Stream() for _ in range(4)]  # test_tilelang_jit_gemm_cython.py:344 multi-stream use
    torch.cuda.synchronize()
    for s in streams:
        with torch.cuda.stream(s):
            outs.append(kern(a, b))
    torch.cuda.synchronize()
    for o in outs[1:]:
        assert torch.equal(o.view(torch.int16), outs[0].view(torch.int16))


UPSTREAM = r'''
import sys, json, torch
import tilelang, tilelang.language as T          # the upstream TileLang wheel of the image (reference lineage)
assert "site-packages" in tilelang.__file__
def matmul(M,N,K,bM,bN,bK,dt,stages,threads):
    @T.prim_func
    def main(A: T.Tensor((M,K),dt), B: T.Tensor((N,K),dt), C: T.Tensor((M,N),dt)):
        with T.Kernel(T.ceildiv(N,bN), T.ceildiv(M,bM), threads=threads) as (bx,by):
            A_s=T.alloc_shared((bM,bK),dt); B_s=T.alloc_shared((bN,bK),dt)
            C_t=T.alloc_tmem([bM,bN],"float"); mbar=T.alloc_barrier(1)
            C_l=T.alloc_fragment((bM,bN),"float"); C_s=T.alloc_shared((bM,bN),dt)
            for k in T.Pipelined(T.ceildiv(K,bK), num_stages=stages):
                T.copy(A[by*bM,k*bK],A_s); T.copy(B[bx*bN,k*bK],B_s)
                T.gemm(A_s,B_s,C_t,False,True,mbar=mbar,clear_accum=k==0)
                T.mbarrier_wait_parity(mbar,k%2)
            T.copy(C_t,C_l); T.copy(C_l,C_s); T.copy(C_s,C[by*bM,bx*bN])
    return main
M,N,K=4096,4096,8192
k=tilelang.compile(matmul(M,N,K,128,256,128,"bfloat16",2,256), out_idx=[2], target="cuda",
    pass_configs={tilelang.PassConfigKey.TL_DISABLE_TMA_LOWER: True, tilelang.PassConfigKey.TL_DISABLE_WARP_SPECIALIZED: True})
torch.manual_seed(0)
a=torch.randn(M,K,device="cuda",dtype=torch.bfloat16); b=torch.randn(N,K,device="cuda",dtype=torch.bfloat16)
c=k(a,b); torch.cuda.synchronize()
lat=k.get_profiler().do_bench()
torch.save(c.cpu(), sys.argv[1])
print("UPSTREAM "+json.dumps({"latency_ms":lat,"tflops":2*M*N*K/lat*1e-9,"version":tilelang.__version__}))
'''


def test_matches_reference_lineage_tcgen05_kernel(tmp_path):
    """Secondary oracle: the reference lineage's own sm_100 tcgen05 kernel, JIT-compiled by
    the upstream TileLang wheel in a separate process, on the same seeded inputs."""
    out = tmp_path / "c_ref.pt"
    env = {k: v for k, v in os.environ.items() if k != "PYTHONPATH"}
    script = tmp_path / "upstream_prog.py"  # upstream's tracer needs inspect.getsource -> a real file
    script.write_text(UPSTREAM)
    try:
        res = subprocess.run([sys.executable, str(script), str(out)], capture_output=True, text=True, timeout=900,
                             env=env, cwd=str(tmp_path))
    except subprocess.TimeoutExpired:
        pytest.skip("upstream TileLang JIT timed out on this box")
    line = [l for l in res.stdout.splitlines() if l.startswith("UPSTREAM ")]
    if res.returncode != 0 or not line:
        pytest.skip("reference GPU kernel unavailable on this box: " + (res.stderr.strip().splitlines() or ["?"])[-1][:300])
    import json
    info = json.loads(line[0][len("UPSTREAM "):])
    c_ref = torch.load(out).cuda()
    M, N, K = 4096, 4096, 8192
    torch.manual_seed(0)
    a = torch.randn(M, K, device="cuda", dtype=torch.bfloat16)
    b = torch.randn(N, K, device="cuda", dtype=torch.bfloat16)
    kern = tilelang.compile(P.tmem_accumulator(M, N, K), out_idx=[2])
    c = kern(a, b)
    # both accumulate in fp32 and round once: they may differ by a bf16 ulp where the K order differs
    diff = (c.float() - c_ref.float()).abs()
    # (tiles of a split tail wave add their K slices' fp32 sums in a different order: with |terms| ~ 1 and sums of
    # magnitude ~90 over K = 8192 that moves an fp32 result by up to ~1e-3 absolute, which near zero is many bf16
    # ulps - hence the absolute floor; elsewhere one bf16 ulp of the larger neighbour across a binade edge)
    ulp = (c_ref.float().abs() * 2.0**-6).clamp_min(4e-3)
    assert (diff <= ulp).all(), f"max diff {diff.max().item()}"
    frac_equal = (c.view(torch.int16) == c_ref.view(torch.int16)).float().mean().item()
    assert frac_equal > 0.95
    lat = kern.get_profiler().do_bench()
    print(f"reference-lineage tcgen05 kernel (tilelang {info['version']}): {info['tflops']:.1f} TFLOP/s; "
          f"this backend: {2 * M * N * K / lat * 1e-9:.1f} TFLOP/s; bit-equal outputs: {frac_equal * 100:.2f}%")
    assert lat < info["latency_ms"]
