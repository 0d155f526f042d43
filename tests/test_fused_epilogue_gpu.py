"""Epilogue vectors fused into the TMEM -> global drain (tlb200_gemm_fused):
C = cast(relu?(((alpha * acc) * row_scale[m]) * col_scale[n] + bias)), every step rounded to fp32 in that order.
Bit-exact against the scalar C oracle (oracle_gemm_fused) on inputs whose products and sums are exact, for all six
output types, every epilogue path (TMA-store slabs, coalesced slabs, direct stores), split-K and the tail split;
tolerance parity on random data against the fp32 torch formula."""
import ctypes

import pytest
import torch

from oracle import gemm_ref as R
from tilelang_mesh_b200 import _abi
from tilelang_mesh_b200 import gemm as G

pytestmark = pytest.mark.gpu
RAW = {1: torch.uint8, 2: torch.int16, 4: torch.int32}


def _exact_inputs(m, n, k, dt, seed=0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    a = torch.randint(-3, 4, (m, k), device="cuda", generator=g).float().to(dt)
    b = torch.randint(-3, 4, (n, k), device="cuda", generator=g).float().to(dt)
    # scales that are powers of two and small-integer biases keep every fp32 step exact, whatever the order
    rs = torch.tensor([0.5, 1.0, 2.0, 0.25], device="cuda")[torch.randint(0, 4, (m,), device="cuda", generator=g)]
    cs = torch.tensor([1.0, 0.5, 4.0, 2.0], device="cuda")[torch.randint(0, 4, (n,), device="cuda", generator=g)]
    return a, b, rs, cs


@pytest.mark.parametrize("out", [torch.float32, torch.bfloat16, torch.float16, torch.float8_e4m3fn, torch.float8_e5m2])
@pytest.mark.parametrize("cfg", [{}, dict(epilogue=1), dict(epilogue=3), dict(cta_group=2, block_n=512), dict(cta_group=1, block_n=64),
                                 dict(split_k=4)])
def test_vectors_bit_exact_vs_c_oracle(out, cfg):
    m, n, k = 300, 528, 264  # n * sizeof(out) is a multiple of 16 for every output type: every epilogue path is legal
    a, b, rs, cs = _exact_inputs(m, n, k, torch.bfloat16)
    bias_n = torch.randint(-4, 5, (n,), device="cuda").float()
    bias_m = torch.randint(-4, 5, (m,), device="cuda").float()
    alpha = 2.0 ** -5
    for kw in (dict(row_scale=rs), dict(col_scale=cs), dict(bias=bias_n), dict(bias=bias_m, bias_per_row=True),
               dict(row_scale=rs, col_scale=cs, bias=bias_n, relu=True)):
        got = G.gemm(a, b, out_dtype=out, alpha=alpha, **kw, **cfg)
        want = R.c_oracle_gemm_fused(a, b, out, alpha=alpha, **kw)
        assert torch.equal(got.cpu().view(RAW[got.element_size()]), want.view(RAW[want.element_size()])), (kw.keys(), cfg)


def test_vectors_int8_operands_and_tail_split():
    # int8 operands accumulate exactly in int32 and leave as float: the vectors apply to the converted value
    m, n, k = 200, 264, 512
    a = torch.randint(-5, 6, (m, k), device="cuda", dtype=torch.int8)
    b = torch.randint(-5, 6, (n, k), device="cuda", dtype=torch.int8)
    rs = torch.full((m,), 0.5, device="cuda")
    bias = torch.arange(n, device="cuda").float()
    got = G.gemm(a, b, out_dtype=torch.float32, row_scale=rs, bias=bias)
    want = (a.float() @ b.float().t()) * 0.5 + bias
    assert torch.equal(got, want)
    # a shape whose automatic plan cuts the last wave along K (reduction kernel applies the vectors)
    m, n, k = 2560, 8192, 8192
    a = torch.randint(-2, 3, (m, k), device="cuda").float().bfloat16()
    b = torch.randint(-2, 3, (n, k), device="cuda").float().bfloat16()
    rs = torch.tensor([0.5, 2.0], device="cuda")[torch.arange(m, device="cuda") % 2]
    bias = (torch.arange(n, device="cuda") % 7).float()
    got = G.gemm(a, b, out_dtype=torch.float32, row_scale=rs, bias=bias, relu=True)
    want = torch.relu((a.float() @ b.float().t()) * rs[:, None] + bias)
    assert torch.equal(got, want)
    assert G.plan(m, n, k)["workspace_bytes"] > 0  # i.e. the plan above did use the tail split


def test_vectors_tolerance_on_random_data():
    m, n, k = 1024, 1536, 2048
    a = torch.randn(m, k, device="cuda").bfloat16()
    b = torch.randn(n, k, device="cuda").bfloat16()
    rs, cs, bias = torch.rand(m, device="cuda") + 0.5, torch.rand(n, device="cuda") + 0.5, torch.randn(n, device="cuda")
    got = G.gemm(a, b, row_scale=rs, col_scale=cs, bias=bias)
    want = ((a.float() @ b.float().t()) * rs[:, None] * cs[None, :] + bias).bfloat16()
    R.assert_allclose(got, want, rtol=1e-2, atol=1e-2)
    # fp8 operands with a per-tensor alpha and per-row scales (the dequantisation form of a row-quantised A)
    a8 = (torch.randn(m, k, device="cuda") * 2).half().to(torch.float8_e4m3fn)
    b8 = (torch.randn(n, k, device="cuda") * 2).half().to(torch.float8_e4m3fn)
    got = G.gemm(a8, b8, out_dtype=torch.bfloat16, alpha=0.25, row_scale=rs)
    want = ((a8.float() @ b8.float().t()) * 0.25 * rs[:, None]).bfloat16()
    R.assert_allclose(got, want, rtol=5e-2, atol=5e-2)


def test_vector_argument_errors_through_the_c_abi():
    lib = _abi.load()
    m = n = k = 256
    a = torch.zeros(m, k, device="cuda", dtype=torch.bfloat16)
    b = torch.zeros(n, k, device="cuda", dtype=torch.bfloat16)
    c = torch.full((m, n), 3.0, device="cuda", dtype=torch.bfloat16)
    v = torch.ones(m, device="cuda")
    desc = G.make_desc(m, n, k, _abi.BFLOAT16, _abi.BFLOAT16, 0, 1, k, k, n, vec_flags=_abi.VEC_BIAS_N | _abi.VEC_BIAS_M)
    vec = _abi.EpilogueVectors(None, None, v.data_ptr())
    assert lib.tlb200_gemm_fused(ctypes.byref(desc), a.data_ptr(), b.data_ptr(), c.data_ptr(), ctypes.byref(vec), None, 0, None) != 0
    assert "per column or per row" in _abi.last_error()
    desc = G.make_desc(m, n, k, _abi.BFLOAT16, _abi.BFLOAT16, 0, 1, k, k, n, vec_flags=_abi.VEC_ROW_SCALE)
    vec = _abi.EpilogueVectors(None, None, None)
    assert lib.tlb200_gemm_fused(ctypes.byref(desc), a.data_ptr(), b.data_ptr(), c.data_ptr(), ctypes.byref(vec), None, 0, None) != 0
    assert "null" in _abi.last_error()
    desc = G.make_desc(m, n, k, _abi.BFLOAT16, _abi.BFLOAT16, 0, 1, k, k, n, vec_flags=_abi.VEC_ROW_SCALE, accumulate=True)
    vec = _abi.EpilogueVectors(v.data_ptr(), None, None)
    assert lib.tlb200_gemm_fused(ctypes.byref(desc), a.data_ptr(), b.data_ptr(), c.data_ptr(), ctypes.byref(vec), None, 0, None) != 0
    desc = G.make_desc(m, n, k, _abi.BFLOAT16, _abi.BFLOAT16, 0, 1, k, k, n, vec_flags=64)
    assert lib.tlb200_gemm_fused(ctypes.byref(desc), a.data_ptr(), b.data_ptr(), c.data_ptr(), ctypes.byref(vec), None, 0, None) != 0
    assert "vec_flags" in _abi.last_error()
    # the plain entry points ignore vec_flags (they are given no vectors)
    desc = G.make_desc(m, n, k, _abi.BFLOAT16, _abi.BFLOAT16, 0, 1, k, k, n, vec_flags=_abi.VEC_ROW_SCALE)
    assert lib.tlb200_gemm(ctypes.byref(desc), a.data_ptr(), b.data_ptr(), c.data_ptr(), None) == 0
    torch.cuda.synchronize()
    assert (c == 0).all()
    with pytest.raises(ValueError, match="row_scale"):
        G.gemm(a, b, row_scale=torch.ones(m + 1, device="cuda"))
