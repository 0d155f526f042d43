"""GPU parity tests (marked gpu): the CUDA path, called through the C ABI (tlb200_gemm via the
ctypes adapter), against the oracle on the same seeded inputs.

Bars (north_star): fp32 accumulate; rel. err <= 1e-2 for bf16/fp16 and <= 5e-2 for fp8 against an
fp32 torch.matmul oracle, judged with the reference's own rule (isclose(rtol, atol=rtol), <= 1 %
mismatches: tilelang/profiler/__init__.py:76-83); integer-valued inputs, whose fp32 sums are exact
in any order, must match the C oracle BIT-EXACTLY.
"""
import ctypes

import pytest
import torch

from oracle import gemm_ref as R
from tilelang_mesh_b200 import _abi
from tilelang_mesh_b200 import gemm as G

pytestmark = pytest.mark.gpu

F8 = (torch.float8_e4m3fn, torch.float8_e5m2)
RAW = {1: torch.uint8, 2: torch.int16, 4: torch.int32}


def rand(shape, dt, kind="randn"):
    if kind == "int":
        return torch.randint(-3, 4, shape, device="cuda").float().to(dt)
    x = torch.randn(shape, device="cuda")
    return x.half().to(dt) if dt in F8 else x.to(dt)  # fp8 supply as example_tilelang_gemm_fp8_sm100.py:111-112


def operands(m, n, k, dt, ta=False, tb=True, kind="randn"):
    return rand((k, m) if ta else (m, k), dt, kind), rand((n, k) if tb else (k, n), dt, kind)


def check_tolerance(c, a, b, ta=False, tb=True, alpha=1.0, relu=False):
    ref = R.matmul_fp32(a, b, ta, tb, alpha, relu)
    tol = 5e-2 if a.dtype in F8 else 1e-2
    R.assert_allclose(c, ref.to(c.dtype), rtol=tol, atol=tol, max_mismatched_ratio=0.01)
    assert R.max_rel_err(c.float(), ref) <= (tol if c.dtype not in F8 else 0.13), "north_star rel. err bound"
    assert R.calc_diff(c.float(), ref) < 1e-3  # fp8 tests' metric (test_tilelang_kernel_fp8_gemm.py:54)


def bit_equal(x, y):
    return torch.equal(x.view(RAW[x.element_size()]), y.view(RAW[y.element_size()]))


# ------------------------------------------------------------------ the reference's own sm_100 grid
# maint/gemm_v2/correctness_evaluation_tcgen05.py:137-190, verbatim shapes: (m, n, 3*k), NT,
# fp16->fp32 for k in {16,32,64,128}; e5m2->fp32 for k in {32,64,128}.
@pytest.mark.parametrize("m", [32, 64, 128, 256])
@pytest.mark.parametrize("n", [64, 128, 256, 512])
@pytest.mark.parametrize("k,dt", [(16, torch.float16), (32, torch.float16), (64, torch.float16), (128, torch.float16),
                                  (32, torch.float8_e5m2), (64, torch.float8_e5m2), (128, torch.float8_e5m2)])
def test_reference_tcgen05_grid(m, n, k, dt):
    a, b = operands(m, n, 3 * k, dt)
    c = G.gemm(a, b, out_dtype=torch.float32)
    ref = R.matmul_fp32(a, b)
    R.assert_allclose(c, ref, rtol=1e-2, atol=1e-2)
    if m * n * k <= 128 * 128 * 64:  # scalar C oracle finishes in well under a second here
        want = R.c_oracle_gemm(a, b, torch.float32)
        assert (c.cpu() - want).abs().max().item() <= 1e-3 * max(1.0, want.abs().max().item())


# ------------------------------------------------------------------ bit-exact parity on exact inputs
@pytest.mark.parametrize("ta,tb", [(False, True), (False, False), (True, True), (True, False)], ids=["nt", "nn", "tn", "tt"])
@pytest.mark.parametrize("dt,out", [(torch.bfloat16, torch.bfloat16), (torch.float16, torch.float32),
                                    (torch.float8_e4m3fn, torch.bfloat16), (torch.float8_e5m2, torch.float16)])
@pytest.mark.parametrize("cg,bn", [(1, 0), (2, 0), (2, 512)])
def test_bit_exact_vs_c_oracle_integer_inputs(ta, tb, dt, out, cg, bn):
    m, n, k = 192, 320, 256
    a, b = operands(m, n, k, dt, ta, tb, "int")
    c = G.gemm(a, b, trans_a=ta, trans_b=tb, out_dtype=out, cta_group=cg, block_n=bn)
    want = R.c_oracle_gemm(a, b, out, ta, tb)
    assert bit_equal(c.cpu(), want)


@pytest.mark.parametrize("epi", [1, 2], ids=["direct", "tma_store"])
@pytest.mark.parametrize("out", [torch.bfloat16, torch.float16, torch.float32, torch.float8_e4m3fn, torch.float8_e5m2])
def test_bit_exact_output_casts(epi, out):
    # half-integer products exercise the round-to-nearest-even of every output type
    m, n, k = 130, 272, 64
    a = (torch.randint(-4, 5, (m, k), device="cuda").float() * 0.5).bfloat16()
    b = (torch.randint(-4, 5, (n, k), device="cuda").float() * 0.25).bfloat16()
    c = G.gemm(a, b, out_dtype=out, alpha=0.125, relu=True, epilogue=epi)
    want = R.c_oracle_gemm(a, b, out, alpha=0.125, relu=True)
    assert bit_equal(c.cpu(), want)


# ------------------------------------------------------------------ ragged / edge shapes
@pytest.mark.parametrize("m,n,k", [(1, 8, 8), (1, 1, 16), (7, 5, 24), (127, 129, 72), (128, 256, 64), (129, 257, 136),
                                   (200, 300, 136), (300, 200, 1032), (513, 1000, 264), (255, 8, 8), (2, 4096, 16)])
@pytest.mark.parametrize("cg", [0, 1, 2])
def test_ragged_shapes_bit_exact(m, n, k, cg):
    a, b = operands(m, n, k, torch.bfloat16, kind="int")
    c = G.gemm(a, b, out_dtype=torch.float32, cta_group=cg)
    want = a.float() @ b.float().t()  # exact: integer-valued inputs
    assert torch.equal(c, want)


@pytest.mark.parametrize("cfg", [dict(), dict(cta_group=3, block_n=128), dict(cta_group=3, block_n=64), dict(cta_group=3, block_n=256, epi_warps=8),
                                 dict(cta_group=2, block_n=64, kgroup=4)], ids=["auto", "h128", "h64", "h256w8", "c2n64g4"])
def test_untouched_outside_of_c_window(cfg):
    # C is a strided window of a larger buffer: nothing outside [m, n] may be written (rows above and below, columns right)
    m, n, k, ldc = 200, 136, 64, 160
    a, b = operands(m, n, k, torch.bfloat16, kind="int")
    for epi in (1, 2):
        if epi == 2 and cfg.get("block_n") == 64 and cfg.get("cta_group") == 3:
            continue  # 32 columns per warp do not fill a 128-byte staging slab of bf16
        big = torch.full((m + 6, ldc), -7.0, device="cuda", dtype=torch.bfloat16)
        c = big[3:m + 3, :n]
        G.gemm(a, b, c, epilogue=epi, **cfg)
        assert torch.equal(c.float(), a.float() @ b.float().t())
        assert (big[:, n:] == -7).all() and (big[m + 3:] == -7).all() and (big[:3] == -7).all()


@pytest.mark.parametrize("dt,n,pad", [(torch.bfloat16, 500, 12), (torch.float16, 1001, 7), (torch.float32, 333, 3),
                                      (torch.float32, 334, 6), (torch.bfloat16, 250, 6), (torch.float8_e4m3fn, 1000, 24)])
def test_padding_right_of_c_untouched_when_row_bytes_not_16_aligned(dt, n, pad):
    # TMA stores clip at the last column only in whole 16-byte units: such rows must take the predicated stores
    m, k = 300, 256
    a, b = operands(m, n, k, torch.bfloat16, kind="int")
    big = torch.full((m, n + pad), -1.0, device="cuda", dtype=torch.float32).to(dt)
    G.gemm(a, b, big[:, :n], alpha=2.0 ** -6)
    assert (big[:, n:].float() == -1).all()
    want = ((a.float() @ b.float().t()) * 2.0 ** -6).to(dt)
    assert torch.equal(big[:, :n].float(), want.float())
    with pytest.raises(RuntimeError, match="row length"):
        G.gemm(a, b, big[:, :n], epilogue=2)


@pytest.mark.parametrize("m,n,k", [(256, 128, 64), (300, 200, 136), (512, 2048, 1024), (1000, 3000, 520), (2048, 2048, 2048)])
@pytest.mark.parametrize("dt", [torch.bfloat16, torch.float8_e4m3fn, torch.float16])
def test_a_multicast_clusters_bit_exact(m, n, k, dt):
    # cfg_cta_group = 4: clusters of two pairs on neighbouring column tiles, A boxes TMA-multicast to both
    a, b = operands(m, n, k, dt, kind="int")
    want = a.float() @ b.float().t()
    for bn in (64, 128, 256):
        if n <= bn:
            continue
        assert G.plan(m, n, k, dt, torch.float32, cta_group=4, block_n=bn)["multicast"] == 2
        c = G.gemm(a, b, out_dtype=torch.float32, cta_group=4, block_n=bn)
        assert torch.equal(c, want), f"block_n={bn}"
    if k >= 1024:  # with K slices through the workspace, and an odd number of column tiles
        c = G.gemm(a, b[:n - 70], out_dtype=torch.float32, cta_group=4, block_n=128, split_k=2)
        assert torch.equal(c, want[:, :n - 70])
    if m % 16 == 0 and dt != torch.float8_e4m3fn:  # an M-contiguous A cannot be split into multicast halves
        with pytest.raises(RuntimeError, match="cfg_cta_group=4"):
            G.gemm(a.t().contiguous(), b, out_dtype=torch.float32, trans_a=True, cta_group=4, block_n=64)


@pytest.mark.parametrize("out", [torch.float32, torch.bfloat16, torch.float16, torch.float8_e4m3fn])
@pytest.mark.parametrize("cfg", [{}, dict(cta_group=1, block_n=64), dict(cta_group=2, block_n=256), dict(cta_group=4, block_n=128),
                                 dict(split_k=4), dict(cta_group=2, block_n=512)])
def test_accumulate_into_c(out, cfg):
    # TLB200_EPI_ACCUMULATE: C = C + alpha * A B^T (examples/gemm_splitk's atomic-add form, done without atomics)
    m, n, k = 384, 520, 2048
    a, b = operands(m, n, k, torch.bfloat16, kind="int")
    c0 = torch.randint(-8, 9, (m, n), device="cuda").float()
    c = c0.to(out).clone()
    alpha = 2.0 ** -7
    G.gemm(a, b, c, alpha=alpha, accumulate=True, **cfg)
    want = ((a.float() @ b.float().t()) * alpha + c0.to(out).float()).to(out)  # exact in fp32, one rounding
    assert torch.equal(c.float(), want.float())
    with pytest.raises(RuntimeError, match="ACCUMULATE"):
        G.gemm(a, b, c, relu=True, accumulate=True)
    with pytest.raises(ValueError, match="accumulate"):
        G.gemm(a, b, accumulate=True)


def test_strided_operands():
    m, n, k = 160, 192, 128
    abig = rand((m, k + 64), torch.bfloat16, "int")
    bbig = rand((n, k + 32), torch.bfloat16, "int")
    a, b = abig[:, :k], bbig[:, 16:16 + k]  # lda = k+64; B window starts 32 bytes in (16 B aligned)
    c = G.gemm(a, b, out_dtype=torch.float32)
    assert torch.equal(c, a.float() @ b.float().t())


def test_empty_extents():
    a = torch.zeros(0, 64, device="cuda", dtype=torch.bfloat16)
    b = torch.zeros(32, 64, device="cuda", dtype=torch.bfloat16)
    assert G.gemm(a, b).shape == (0, 32)
    a = torch.zeros(16, 0, device="cuda", dtype=torch.bfloat16)
    b = torch.zeros(32, 0, device="cuda", dtype=torch.bfloat16)
    c = torch.full((16, 32), 5.0, device="cuda", dtype=torch.bfloat16)
    G.gemm(a, b, c)  # empty contraction == T.clear(C)
    assert (c == 0).all()


def test_k_tail_is_zero_filled_not_garbage():
    # K = 72 is not a multiple of the 64-element K block: the TMA box must zero-fill the tail
    m, n, k = 128, 128, 72
    a, b = operands(m, n, k, torch.bfloat16, kind="int")
    assert torch.equal(G.gemm(a, b, out_dtype=torch.float32), a.float() @ b.float().t())


# ------------------------------------------------------------------ tolerance parity on random data
@pytest.mark.parametrize("dt,out", [(torch.bfloat16, torch.bfloat16), (torch.float16, torch.float16),
                                    (torch.float8_e4m3fn, torch.bfloat16), (torch.float8_e5m2, torch.bfloat16)])
@pytest.mark.parametrize("ta,tb", [(False, True), (False, False), (True, True), (True, False)], ids=["nt", "nn", "tn", "tt"])
def test_random_tolerance_1k(dt, out, ta, tb):
    m, n, k = 1024, 1024, 1024  # testing/python/kernel/test_tilelang_kernel_fp8_gemm.py uses 1024^3
    a, b = operands(m, n, k, dt, ta, tb)
    c = G.gemm(a, b, trans_a=ta, trans_b=tb, out_dtype=out)
    check_tolerance(c, a, b, ta, tb)


def test_reference_kernel_gemm_shape():
    # testing/python/kernel/test_tilelang_kernel_gemm.py:107-300 run 512x1024x768 in many dtypes
    for dt, out in ((torch.float16, torch.float16), (torch.bfloat16, torch.bfloat16), (torch.float16, torch.float32)):
        a, b = operands(512, 1024, 768, dt)
        check_tolerance(G.gemm(a, b, out_dtype=out), a, b)


def test_fp8_per_tensor_scales():
    # C3: e4m3 operands, per-tensor scales folded into alpha = scale_a * scale_b, bf16 output
    m = n = k = 2048
    af, bf = torch.randn(m, k, device="cuda") * 3.0, torch.randn(n, k, device="cuda") * 0.02
    sa, sb = af.abs().max().item() / 448.0, bf.abs().max().item() / 448.0
    a, b = (af / sa).to(torch.float8_e4m3fn), (bf / sb).to(torch.float8_e4m3fn)
    c = G.gemm(a, b, out_dtype=torch.bfloat16, alpha=sa * sb)
    check_tolerance(c, a, b, alpha=sa * sb)
    # and against the UNQUANTISED product the scales are meant to recover (looser: fp8 quantisation)
    assert R.calc_diff(c.float(), af @ bf.t()) < 5e-3


@pytest.mark.parametrize("dt,out", [(torch.bfloat16, torch.bfloat16), (torch.float8_e4m3fn, torch.bfloat16)], ids=["C2_bf16", "C3_fp8"])
def test_headline_8192_cubed(dt, out):
    m = n = k = 8192
    a, b = operands(m, n, k, dt)
    c = G.gemm(a, b, out_dtype=out)
    check_tolerance(c, a, b)
    # size-independent property: C @ 1 == A @ (B^T @ 1) — a checksum of every output element
    rows = c.double().sum(dim=1)
    want = a.double() @ b.double().sum(dim=0)
    scale = (a.double().abs() @ b.double().abs().sum(dim=0)).clamp_min(1.0)
    assert ((rows - want).abs() / scale).max().item() < (2e-2 if dt in F8 else 4e-3)
    # determinism + M-shard invariance: a row block computed alone is bit-identical to the same rows of the full run
    c2 = G.gemm(a, b, out_dtype=out)
    assert bit_equal(c, c2)
    part = G.gemm(a[4096:5120], b, out_dtype=out)
    assert bit_equal(part, c[4096:5120])


@pytest.mark.parametrize("m,n,k", [(16384, 14336, 4096), (16384, 4096, 14336)], ids=["C4_mlp_up", "C4_mlp_down"])
def test_llama_mlp_shapes(m, n, k):
    a, b = operands(m, n, k, torch.bfloat16)
    c = G.gemm(a, b)
    check_tolerance(c, a, b)


@pytest.mark.parametrize("cfg", [dict(cta_group=1, block_n=128), dict(cta_group=2, block_n=64), dict(cta_group=2, block_n=192),
                                 dict(cta_group=1, block_n=96, stages=3), dict(cta_group=2, block_n=32), dict(cta_group=2, block_n=512),
                                 dict(cta_group=2, block_n=512, epi_warps=4, epilogue=1), dict(cta_group=1, block_n=384, stages=2), dict(cta_group=2, block_n=448), dict(cta_group=2, raster=1), dict(cta_group=2, raster=64, stages=2)])
def test_tuning_knobs_do_not_change_results(cfg):
    a, b = operands(1000, 1200, 520, torch.bfloat16)
    base = G.gemm(a, b, cta_group=1, block_n=256, epilogue=1)
    assert bit_equal(G.gemm(a, b, **cfg), base)  # same K order per element => identical bits


@pytest.mark.parametrize("kgroup", [1, 2, 4])
@pytest.mark.parametrize("cfg", [dict(cta_group=1, block_n=64), dict(cta_group=2, block_n=64), dict(cta_group=2, block_n=128),
                                 dict(cta_group=1, block_n=32, stages=5), dict(cta_group=4, block_n=64), dict(cta_group=2, block_n=96)])
@pytest.mark.parametrize("k", [64, 192, 448, 1160])  # 1, 3, 7 and 19 K blocks (the last one partial): groups end ragged
def test_stage_groups_do_not_change_results(kgroup, cfg, k):
    """K blocks that share a barrier pair (cfg_kgroup) change the hand-shake, not the arithmetic: same bits as one K
    block per barrier, for K-block counts that are not multiples of the group and more tiles than CTAs."""
    a, b = operands(700, 8200 if cfg["block_n"] <= 64 else 9300, k, torch.bfloat16)
    base = G.gemm(a, b, cta_group=1, block_n=256, epilogue=1)
    assert bit_equal(G.gemm(a, b, kgroup=kgroup, **cfg), base)


@pytest.mark.parametrize("bn", [64, 128, 256])
@pytest.mark.parametrize("dt,out", [(torch.bfloat16, torch.bfloat16), (torch.float8_e4m3fn, torch.float32), (torch.float16, torch.float16),
                                    (torch.int8, torch.int32)])
@pytest.mark.parametrize("m,n,k", [(65, 128, 64), (128, 256, 128), (300, 200, 136), (1000, 1200, 520), (512, 2048, 1160)])
def test_pair_tiles_of_128_rows_bit_exact(bn, dt, out, m, n, k):
    """cfg_cta_group = 3: cta_group::2 MMAs with M = 128, 64 rows per CTA, the accumulator laid out as two column halves
    over the 128 TMEM lanes.  Integer-valued operands: every plan must give the oracle's bits."""
    a, b = operands(m, n, k, dt, kind="int")
    c = G.gemm(a, b, out_dtype=out, cta_group=3, block_n=bn)
    assert bit_equal(c, G.gemm(a, b, out_dtype=out, cta_group=1, block_n=256, epilogue=1))
    for extra in (dict(epilogue=1), dict(epi_warps=8) if bn >= 128 else dict(stages=3), dict(split_k=2), dict(kgroup=1)):
        if "split_k" in extra and dt == torch.int8:
            continue
        assert bit_equal(G.gemm(a, b, out_dtype=out, cta_group=3, block_n=bn, **extra), c), extra


def test_pair_tiles_of_128_rows_argument_errors():
    a, b = operands(64, 128, 64, torch.bfloat16)
    with pytest.raises(RuntimeError, match="cfg_cta_group=3"):
        G.gemm(a, b, cta_group=3, block_n=128)  # needs more than 64 rows
    a, b = operands(256, 256, 64, torch.bfloat16)
    with pytest.raises(RuntimeError, match="multiple of 64"):
        G.gemm(a, b, cta_group=3, block_n=96)
    with pytest.raises(RuntimeError, match="K-contiguous A"):
        G.gemm(a.t().contiguous(), b, trans_a=True, cta_group=3, block_n=128)


# ------------------------------------------------------------------ tf32 and int8 operands
# testing/python/kernel/test_tilelang_kernel_gemm.py:156-247: f32f32f32 nn/nt/tn and i8i8i32 nt/tn at 512x1024x768
@pytest.mark.parametrize("ta,tb", [(False, False), (False, True), (True, False), (True, True)], ids=["nn", "nt", "tn", "tt"])
@pytest.mark.parametrize("cg", [1, 2])
def test_float32_operands_run_as_tf32(ta, tb, cg):
    m, n, k = 512, 1024, 768
    a = torch.randn((k, m) if ta else (m, k), device="cuda")
    b = torch.randn((n, k) if tb else (k, n), device="cuda")
    c = G.gemm(a, b, trans_a=ta, trans_b=tb, cta_group=cg)
    assert c.dtype == torch.float32
    R.assert_allclose(c, R.matmul_fp32(a, b, ta, tb), rtol=1e-2, atol=1e-2)
    ai = torch.randint(-3, 4, a.shape, device="cuda").float()
    bi = torch.randint(-3, 4, b.shape, device="cuda").float()
    assert torch.equal(G.gemm(ai, bi, trans_a=ta, trans_b=tb, cta_group=cg), R.matmul_fp32(ai, bi, ta, tb))  # exact in tf32


@pytest.mark.parametrize("ta,tb", [(False, True), (True, False), (False, False)], ids=["nt", "tn", "nn"])
@pytest.mark.parametrize("out", [torch.int32, torch.float32, torch.bfloat16])
def test_int8_operands_int32_accumulate(ta, tb, out):
    m, n, k = 512, 1024, 768
    a = torch.randint(-128, 128, (k, m) if ta else (m, k), device="cuda", dtype=torch.int8)
    b = torch.randint(-128, 128, (n, k) if tb else (k, n), device="cuda", dtype=torch.int8)
    want = (a.cpu().int().t() if ta else a.cpu().int()) @ (b.cpu().int().t() if tb else b.cpu().int())
    c = G.gemm(a, b, trans_a=ta, trans_b=tb, out_dtype=out)
    if out == torch.int32:
        assert torch.equal(c.cpu(), want)  # bit-exact integer arithmetic
        assert torch.equal(G.gemm(a, b, trans_a=ta, trans_b=tb, out_dtype=out, relu=True).cpu(), want.clamp_min(0))
    else:
        assert torch.equal(c.cpu(), want.float().to(out))
    with pytest.raises(RuntimeError, match="alpha"):
        G.gemm(a, b, trans_a=ta, trans_b=tb, out_dtype=torch.int32, alpha=0.5)


# ------------------------------------------------------------------ operands TMA cannot address directly
# testing/python/kernel/test_tilelang_kernel_gemm.py:250-300 (test_pad_*): odd extents, NN layout
@pytest.mark.parametrize("m,n,k,out", [(512 - 8, 1024 - 32, 768 - 24, torch.float16), (512 - 9, 1024 - 7, 768 - 5, torch.float16),
                                       (512 + 19, 1024 + 17, 768 + 15, torch.float32)], ids=["pad_aligned", "pad", "pad_f32"])
def test_reference_pad_cases_nn(m, n, k, out):
    a, b = operands(m, n, k, torch.float16, False, False)
    c = G.gemm(a, b, trans_a=False, trans_b=False, out_dtype=out)
    R.assert_allclose(c, R.matmul_fp32(a, b, False, False).to(out), rtol=1e-2, atol=1e-2)
    a, b = operands(m, n, k, torch.float16, False, False, "int")
    assert torch.equal(G.gemm(a, b, trans_a=False, trans_b=False, out_dtype=torch.float32), a.float() @ b.float())


@pytest.mark.parametrize("ta,tb", [(False, True), (True, False), (True, True)], ids=["nt", "tt", "tn"])
@pytest.mark.parametrize("dt", [torch.bfloat16, torch.float8_e4m3fn])
def test_unaligned_pitches_and_bases(ta, tb, dt):
    m, n, k = 203, 333, 1001
    a, b = operands(m, n, k, dt, ta, tb, "int")
    want = R.matmul_fp32(a, b, ta, tb)
    assert torch.equal(G.gemm(a, b, trans_a=ta, trans_b=tb, out_dtype=torch.float32), want)
    # views whose base is not 16-byte aligned
    big_a = rand((a.shape[0] + 1, a.shape[1] + 3), dt, "int")
    av = big_a[1:, 3:]
    assert torch.equal(G.gemm(av, b, trans_a=ta, trans_b=tb, out_dtype=torch.float32), R.matmul_fp32(av, b, ta, tb))


# ------------------------------------------------------------------ grouped contraction (MoE expert GEMMs)
def torch_gmm(a, b, rows, trans_b):
    # the reference's checker: examples/grouped_gemm/example_grouped_gemm_fwd.py:8-38
    out, start = [], 0
    for g, r in enumerate(rows):
        bg = b[g].float().t() if trans_b else b[g].float()
        out.append(a[start:start + r].float() @ bg)
        start += r
    return torch.cat(out) if out else a.new_zeros((0, b.shape[1] if trans_b else b.shape[2])).float()


@pytest.mark.parametrize("rows", [[64, 128, 512, 33], [300, 0, 1, 700, 256], [4096, 4096], [7] * 40, [129, 255, 257, 511],
                                  [5, 0, 130] * 85 + [1], [64] * 128])  # the last two: 256 and 128 groups (expert counts of MoE layers)
@pytest.mark.parametrize("trans_b", [False, True], ids=["b_kn", "b_nk"])
@pytest.mark.parametrize("cfg", [dict(), dict(cta_group=1, block_n=128), dict(cta_group=2, block_n=256), dict(cta_group=3, block_n=128),
                                 dict(cta_group=3, block_n=256)], ids=["auto", "c1", "c2", "h128", "h256"])
def test_grouped_gemm(rows, trans_b, cfg):
    k, n = 512, 768
    a = rand((sum(rows), k), torch.bfloat16, "int")
    b = rand((len(rows), n, k) if trans_b else (len(rows), k, n), torch.bfloat16, "int")
    c = torch.full((sum(rows), n), 7.0, device="cuda", dtype=torch.float32)
    G.gemm_grouped(a, b, rows, c, trans_b=trans_b, **cfg)
    assert torch.equal(c, torch_gmm(a, b, rows, trans_b))  # exact inputs: bit-exact


def test_grouped_gemm_random_fp16_and_errors():
    rows = [1000, 24, 3000, 72]  # example_grouped_gemm_fwd.py-style uneven batch sizes
    k, n = 4096, 2048
    a = rand((sum(rows), k), torch.float16)
    b = rand((len(rows), k, n), torch.float16)
    c = G.gemm_grouped(a, b, rows)
    R.assert_allclose(c, torch_gmm(a, b, rows, False).half(), 1e-2, 1e-2)
    with pytest.raises(ValueError, match="sum to"):
        G.gemm_grouped(a, b, [1, 2, 3, 4])
    with pytest.raises(RuntimeError, match="n_groups"):
        G.gemm_grouped(rand((257, 64), torch.float16), rand((257, 64, 64), torch.float16), [1] * 257)


# ------------------------------------------------------------------ split-K (scratch workspace)
@pytest.mark.parametrize("m,n,k", [(128, 4096, 14336), (256, 256, 16384), (100, 520, 4112), (384, 1024, 8192)])
@pytest.mark.parametrize("dt,out", [(torch.bfloat16, torch.bfloat16), (torch.float8_e4m3fn, torch.float32)])
def test_split_k_matches_unsplit(m, n, k, dt, out):
    assert G.plan(m, n, k, dt, out, split_k=4)["split_k"] > 1
    a, b = operands(m, n, k, dt, kind="int")
    c = G.gemm(a, b, out_dtype=out, alpha=0.5, relu=True, split_k=4)   # split-K through tlb200_gemm_ex
    base = G.gemm(a, b, out_dtype=out, alpha=0.5, relu=True, split_k=1)
    assert bit_equal(c, base)                                        # exact inputs: any summation order gives the same bits
    assert bit_equal(c.cpu(), torch.relu(0.5 * (a.float() @ b.float().t())).to(out).cpu())
    a, b = operands(m, n, k, dt)
    check_tolerance(G.gemm(a, b, out_dtype=torch.bfloat16), a, b)
    for s in (2, 5):
        check_tolerance(G.gemm(a, b, out_dtype=torch.bfloat16, split_k=s, cta_group=1, block_n=128), a, b)


def test_split_k_workspace_contract():
    m, n, k = 128, 1024, 8192
    a, b = operands(m, n, k, torch.bfloat16, kind="int")
    c = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
    lib = _abi.load()
    d = G.make_desc(m, n, k, _abi.BFLOAT16, _abi.BFLOAT16, 0, 1, k, k, n, split_k=4)
    need = lib.tlb200_gemm_workspace_bytes(ctypes.byref(d))
    assert need == 4 * 128 * 1024 * 4
    ws = torch.empty(need, dtype=torch.uint8, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    n0 = lib.tlb200_launch_count()
    assert lib.tlb200_gemm_ex(ctypes.byref(d), a.data_ptr(), b.data_ptr(), c.data_ptr(), ws.data_ptr(), need, st) == 0
    assert lib.tlb200_launch_count() == n0 + 2  # GEMM + reduce
    assert torch.equal(c.float(), (a.float() @ b.float().t()).bfloat16().float())
    # explicit split with too small a workspace is an error; with no workspace through tlb200_gemm too
    assert lib.tlb200_gemm_ex(ctypes.byref(d), a.data_ptr(), b.data_ptr(), c.data_ptr(), ws.data_ptr(), need - 256, st) != 0
    assert "workspace" in _abi.last_error()
    assert lib.tlb200_gemm(ctypes.byref(d), a.data_ptr(), b.data_ptr(), c.data_ptr(), st) != 0
    # automatic split silently runs unsplit when the workspace is too small
    d0 = G.make_desc(m, n, k, _abi.BFLOAT16, _abi.BFLOAT16, 0, 1, k, k, n)
    n0 = lib.tlb200_launch_count()
    assert lib.tlb200_gemm_ex(ctypes.byref(d0), a.data_ptr(), b.data_ptr(), c.data_ptr(), ws.data_ptr(), 1024, st) == 0
    assert lib.tlb200_launch_count() == n0 + 1


# ------------------------------------------------------------------ boundary behaviour
def test_argument_errors():
    a, b = operands(64, 64, 64, torch.bfloat16)
    with pytest.raises(ValueError, match="K shape check"):
        G.gemm(a, b[:, :32])
    with pytest.raises(ValueError, match="is torch.float16"):
        G.gemm(a, b.half())
    with pytest.raises(ValueError, match="CUDA device"):
        G.gemm(a.cpu(), b.cpu())
    with pytest.raises(ValueError, match="expected"):
        G.gemm(a, b, torch.empty(64, 32, device="cuda", dtype=torch.bfloat16))
    with pytest.raises(ValueError, match="contiguous"):
        G.gemm(a.t(), b)
    lib = _abi.load()
    a100, b100 = operands(64, 64, 100, torch.bfloat16)  # 200-byte rows: not addressable by TMA without a workspace
    d100 = G.make_desc(64, 64, 100, _abi.BFLOAT16, _abi.BFLOAT16, 0, 1, 100, 100, 64)
    assert lib.tlb200_gemm(ctypes.byref(d100), a100.data_ptr(), b100.data_ptr(), a.data_ptr(), None) != 0
    assert "multiple of 16 bytes" in _abi.last_error()
    d = G.make_desc(64, 64, 64, _abi.BFLOAT16, _abi.BFLOAT16, 0, 1, 64, 64, 64)
    assert lib.tlb200_gemm(ctypes.byref(d), None, b.data_ptr(), a.data_ptr(), None) != 0
    assert "null tensor pointer" in _abi.last_error()


def test_launches_on_callers_stream_and_counts():
    a, b = operands(512, 512, 512, torch.bfloat16)
    lib = _abi.load()
    n0 = lib.tlb200_launch_count()
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        c1 = G.gemm(a, b)
    s.synchronize()
    c0 = G.gemm(a, b)
    torch.cuda.synchronize()
    assert bit_equal(c0, c1)
    assert lib.tlb200_launch_count() == n0 + 2


@pytest.mark.parametrize("m,n,k", [(512, 384, 256), (4100, 640, 320), (4100, 2600, 320), (2048, 9000, 64)],
                         ids=["one_panel", "row_panels", "a_and_b_panels_alternate", "many_b_panels"])
def test_host_buffer_entry_point(m, n, k):
    a = torch.randint(-3, 4, (m, k)).float().bfloat16().pin_memory()
    b = torch.randint(-3, 4, (n, k)).float().bfloat16().pin_memory()
    c = torch.empty(m, n, dtype=torch.bfloat16).pin_memory()
    lib = _abi.load()
    G._ensure_init(0)
    d = G.make_desc(m, n, k, _abi.BFLOAT16, _abi.BFLOAT16, 0, 1, k, k, n)
    need = lib.tlb200_gemm_host_workspace_bytes(ctypes.byref(d))
    ws = torch.empty(need, dtype=torch.uint8, device="cuda")
    rc = lib.tlb200_gemm_host(ctypes.byref(d), a.data_ptr(), b.data_ptr(), c.data_ptr(), ws.data_ptr(), need,
                              torch.cuda.current_stream().cuda_stream)
    assert rc == 0, _abi.last_error()
    assert torch.equal(c, (a.float() @ b.float().t()).bfloat16())
    assert lib.tlb200_gemm_host(ctypes.byref(d), a.data_ptr(), b.data_ptr(), c.data_ptr(), ws.data_ptr(), need - 1, None) != 0


def test_cuda_graph_capture_and_replay():
    # tlb200_gemm encodes its tensor maps on the host and makes one launch on the given stream: it can be captured.
    # A launch-bound chain of small contractions replayed from a graph must give the same bits as eager calls.
    a, b = operands(256, 256, 256, torch.bfloat16, kind="int")
    w = [rand((256, 256), torch.bfloat16, "int") for _ in range(4)]
    x0 = torch.empty(256, 256, device="cuda", dtype=torch.bfloat16)
    xs = [torch.empty_like(x0) for _ in range(4)]

    def chain():
        G.gemm(a, b, x0, alpha=2.0 ** -8)
        src = x0
        for wi, xi in zip(w, xs):
            G.gemm(src, wi, xi, alpha=2.0 ** -8)
            src = xi

    chain()
    torch.cuda.synchronize()
    eager = [t.clone() for t in xs]
    for t in xs:
        t.zero_()
    g = torch.cuda.CUDAGraph()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        chain()  # warm-up on the capture stream (descriptor cache, device init)
        with torch.cuda.graph(g, stream=side):
            chain()
    torch.cuda.current_stream().wait_stream(side)
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    for got, want in zip(xs, eager):
        assert bit_equal(got, want)


@pytest.mark.parametrize("dt,out", [(torch.bfloat16, torch.float32), (torch.bfloat16, torch.bfloat16), (torch.float8_e4m3fn, torch.float16),
                                    (torch.float16, torch.float8_e4m3fn)])
@pytest.mark.parametrize("cfg", [dict(cta_group=2, block_n=128), dict(cta_group=1, block_n=256), dict(cta_group=2, block_n=256)])
def test_tail_split_bit_exact(dt, out, cfg):
    # cfg_split_k = -1: the tiles of the last, partial wave are cut along K over all clusters and summed by a
    # second kernel.  Same bits as the ordinary schedule on exactly representable inputs, ragged edges included.
    m, n, k = 2000, 2504, 2048
    a, b = operands(m, n, k, dt, kind="int")
    plan = G.plan(m, n, k, dt, out, split_k=-1, **cfg)
    assert plan["workspace_bytes"] > 0 and plan["split_k"] == 1, "the shape must leave a partial last wave"
    want = ((a.float() @ b.float().t()) * 2.0 ** -5).to(out)
    big = torch.full((m + 2, n + 8), -1.0, device="cuda").to(out)
    G.gemm(a, b, big[:m, :n], alpha=2.0 ** -5, split_k=-1, **cfg)
    assert torch.equal(big[:m, :n].float(), want.float())
    assert (big[m:].float() == -1).all() and (big[:, n:].float() == -1).all()
    assert bit_equal(G.gemm(a, b, out_dtype=out, alpha=2.0 ** -5, split_k=1, **cfg), want)  # switched off: same result


def test_tail_split_epilogues_and_small_workspace():
    m, n, k = 2048, 2560, 2048
    a, b = operands(m, n, k, torch.bfloat16, kind="int")
    cfg = dict(cta_group=2, block_n=128, split_k=-1)
    ref = a.float() @ b.float().t()
    assert torch.equal(G.gemm(a, b, out_dtype=torch.float32, relu=True, **cfg), torch.relu(ref))
    c0 = torch.randint(-8, 9, (m, n), device="cuda").float()
    c = c0.clone()
    G.gemm(a, b, c, accumulate=True, **cfg)
    assert torch.equal(c, ref + c0)
    # through the C ABI with a workspace that lacks the tail part: the ordinary schedule runs, same result
    lib = _abi.load()
    d = G.make_desc(m, n, k, _abi.BFLOAT16, _abi.FLOAT32, 0, 1, k, k, n, **cfg)
    need = lib.tlb200_gemm_workspace_bytes(ctypes.byref(d))
    assert need > 0
    out = torch.empty(m, n, device="cuda")
    small = torch.empty(1024, dtype=torch.uint8, device="cuda")
    n0 = lib.tlb200_launch_count()
    _abi.check(lib.tlb200_gemm_ex(ctypes.byref(d), a.data_ptr(), b.data_ptr(), out.data_ptr(), small.data_ptr(), small.numel(),
                                  torch.cuda.current_stream().cuda_stream), "tlb200_gemm_ex")
    assert lib.tlb200_launch_count() - n0 == 1 and torch.equal(out, ref)
    ws = torch.empty(need, dtype=torch.uint8, device="cuda")
    out.zero_()
    n0 = lib.tlb200_launch_count()
    _abi.check(lib.tlb200_gemm_ex(ctypes.byref(d), a.data_ptr(), b.data_ptr(), out.data_ptr(), ws.data_ptr(), need,
                                  torch.cuda.current_stream().cuda_stream), "tlb200_gemm_ex")
    assert lib.tlb200_launch_count() - n0 == 2 and torch.equal(out, ref)  # GEMM + tail reduction


def test_host_buffer_entry_point_repitches_unaligned_operands():
    # K = 763 fp16: a host pitch of 1526 bytes is not TMA-addressable; the copy to the device re-pitches the rows
    # (the reference's test_pad_* shapes, through the host-buffer form of the call)
    m, n, k = 1100, 520, 763
    a = torch.randint(-3, 4, (m, k)).half().pin_memory()
    b = torch.randint(-3, 4, (k, n + 3)).half().pin_memory()[:, :n]  # B given as [K, N] with an odd pitch too
    c = torch.empty(m, n, dtype=torch.float32).pin_memory()
    lib = _abi.load()
    G._ensure_init(0)
    d = G.make_desc(m, n, k, _abi.FLOAT16, _abi.FLOAT32, 0, 0, k, n + 3, n)
    need = lib.tlb200_gemm_host_workspace_bytes(ctypes.byref(d))
    ws = torch.empty(need, dtype=torch.uint8, device="cuda")
    rc = lib.tlb200_gemm_host(ctypes.byref(d), a.data_ptr(), b.data_ptr(), c.data_ptr(), ws.data_ptr(), need,
                              torch.cuda.current_stream().cuda_stream)
    assert rc == 0, _abi.last_error()
    assert torch.equal(c, a.float() @ b.float())
