"""CPU tests: the oracle against the golden fixtures produced by the reference lineage.

These pin oracle/ (the checker) before any GPU parity claim is made with it.
"""
import json
import os
import re

import numpy as np
import pytest
import torch

from oracle import descriptors as D
from oracle import gemm_ref as R

GOLD = os.path.join(os.path.dirname(__file__), "golden")


@pytest.fixture(scope="module")
def lowering():
    with open(os.path.join(GOLD, "tcgen05_lowering.json")) as f:
        return json.load(f)["records"]


def test_golden_covers_supported_and_unsupported(lowering):
    sup = [r for r in lowering if "instr_desc" in r]
    assert len(sup) >= 8
    # the headline example program of the reference (128x256x128 bf16): SURVEY fact a6
    head = [r for r in sup if (r["block_M"], r["block_N"], r["block_K"], r["in_dtype"]) == (128, 256, 128, "bfloat16")]
    assert head and head[0]["instr_desc"] == 138413200 == 0x08400490


def test_instr_desc_matches_reference_lowering(lowering):
    for r in lowering:
        if "instr_desc" not in r:
            continue
        meta = D.tcgen5_mma_meta(r["block_M"], r["block_N"], r["block_K"], r["in_dtype"], "float32")
        assert meta is not None, r
        am, an, ak, ws, two = meta
        assert ws == r["ws"]
        got = D.tcgen5_instr_desc(am, an, ak, r["in_dtype"], "float32", True, True)
        assert got == r["instr_desc"], (r, hex(got))
        assert D.tmem_columns(r["block_N"]) == r["tmem_cols"]


def test_smem_desc_params_match_reference_lowering(lowering):
    for r in lowering:
        if "desc_a" not in r:
            continue
        eb = 2 if r["in_dtype"] in ("bfloat16", "float16") else 1
        lbo, sbo, swz, offs = D.k_major_desc_params(r["block_K"], eb)
        for which, rows in (("desc_a", r["block_M"]), ("desc_b", r["block_N"])):
            g = r[which]
            assert (g["lbo16"], g["sbo16"], g["swizzle"], g["base_offset"], g["lbo_mode"]) == (lbo, sbo, swz, 0, 0)
        assert r["k_atoms"] == len(offs(r["block_M"]))
        # evaluate the K-offset expression the lineage generated, for every ki
        for expr, rows in zip(r["k_offset_exprs"], (r["block_M"], r["block_N"])):
            expr = expr.strip()
            expr = expr + ")" * (expr.count("(") - expr.count(")"))
            for ki, want in enumerate(offs(rows)):
                assert eval(expr, {"ki": ki}) == want, (r, expr, ki)


def test_smem_desc_bits():
    d = D.tcgen05_smem_desc(0x12340, 1, 64, 0, False, 2)
    assert d & 0x3FFF == 0x1234
    assert (d >> 16) & 0x3FFF == 1 and (d >> 32) & 0x3FFF == 64
    assert (d >> 46) & 3 == 1 and (d >> 61) == 2


def test_full_bank_swizzle_is_tma_128b_pattern():
    # The 128B TMA/UMMA swizzle XORs the 16-byte chunk index with (row % 8); the reference's layout
    # function must describe exactly that (so the kernel's TMA boxes and descriptors agree with it).
    for i in range(16):
        for j in range(0, 128, 8):  # bf16: 8 elements per 16 B
            tc, ts, idx = D.full_bank_swizzle_index(i, j, 16)
            chunk = (j // 8) % 8
            assert tc == j // 64 and ts == i // 8
            assert idx == ((chunk ^ (i % 8)) + (i % 8) * 8) * 8


def test_c_oracle_bit_exact_vs_reference_cpu_target_golden():
    g = np.load(os.path.join(GOLD, "cpu_c_target_matmul.npz"))
    A = torch.from_numpy(g["A"]).view(torch.float16)
    B = torch.from_numpy(g["B"]).view(torch.float16)
    C = torch.from_numpy(g["C"]).view(torch.float16)
    got = R.c_oracle_gemm(A, B, torch.float16, trans_a=False, trans_b=False, product_mode=1)
    assert torch.equal(got.view(torch.int16), C.view(torch.int16))
    # and the reference's own acceptance rule for that program holds against torch
    # (test_tilelang_cpu_gemm.py:115: atol=rtol=1e-2, 5 % mismatches allowed)
    R.assert_allclose(C, torch.matmul(A.float(), B.float()).half(), 1e-2, 1e-2, 0.05)


@pytest.mark.parametrize("dt", [torch.float16, torch.bfloat16, torch.float8_e4m3fn, torch.float8_e5m2])
def test_c_oracle_codec_matches_torch(dt):
    lib = R._c()
    x = torch.randn(2048) * torch.logspace(-8, 2, 2048)
    x = torch.cat([x, torch.tensor([0.0, -0.0, 1.0, -1.0, 0.1, 3.0, 6e-8, 2.0**-9, 2.0**-10, 1.5 * 2.0**-10, 240.0, 447.0])])
    want = x.to(dt)
    got = torch.empty_like(want)
    for i in range(x.numel()):
        lib.oracle_encode(R.DT[dt], got.data_ptr(), i, float(x[i]))
    raw = torch.uint8 if dt.itemsize == 1 else torch.int16
    assert torch.equal(got.view(raw), want.view(raw))
    dec = torch.tensor([lib.oracle_decode(R.DT[dt], want.data_ptr(), i) for i in range(want.numel())])
    assert torch.equal(dec, want.float())


def test_c_oracle_fp8_saturates_like_cvt_satfinite():
    lib = R._c()
    buf = torch.zeros(4, dtype=torch.uint8)
    for i, v in enumerate([1e9, -1e9, 465.0, 449.0]):
        lib.oracle_encode(R.DT[torch.float8_e4m3fn], buf.data_ptr(), i, v)
    assert buf.view(torch.float8_e4m3fn).float().tolist() == [448.0, -448.0, 448.0, 448.0]


@pytest.mark.parametrize("ab,out", [(torch.bfloat16, torch.bfloat16), (torch.float16, torch.float32),
                                    (torch.float8_e4m3fn, torch.bfloat16)])
@pytest.mark.parametrize("ta,tb", [(False, True), (False, False), (True, True), (True, False)])
def test_c_oracle_tensor_mode_vs_fp32_matmul(ab, out, ta, tb):
    m, n, k = 24, 40, 32
    a = torch.randn((k, m) if ta else (m, k)).to(ab)
    b = torch.randn((n, k) if tb else (k, n)).to(ab)
    ref = R.matmul_fp32(a, b, ta, tb, alpha=0.5, relu=True)
    got = R.c_oracle_gemm(a, b, out, ta, tb, alpha=0.5, relu=True)
    assert R.max_rel_err(got.float(), ref) <= (1e-2 if out != torch.float32 else 1e-5)


def test_c_oracle_integer_inputs_exact():
    a = torch.randint(-3, 4, (16, 64)).float().bfloat16()
    b = torch.randint(-3, 4, (24, 64)).float().bfloat16()
    got = R.c_oracle_gemm(a, b, torch.float32)
    assert torch.equal(got, a.float() @ b.float().t())


def test_c_oracle_blockscaled_known_answer_and_identity():
    # examples/deepseek_deepgemm/example_deepgemm_fp8_2xAcc.py:43-83 restated.  Known answer by hand: with
    # A = B = all ones, block kb contributes 128 * sa[m,kb] * sb[n//128,kb] (last block: K - 128*kb ones).
    m, n, k = 3, 130, 300
    a = torch.ones(m, k).to(torch.float8_e4m3fn)
    b = torch.ones(n, k).to(torch.float8_e4m3fn)
    sa = torch.tensor([[1.0, 2.0, 4.0], [0.5, 0.5, 0.5], [0.0, 1.0, 0.0]])
    sb = torch.tensor([[1.0, 1.0, 1.0], [2.0, 0.0, 0.25]])
    c = R.c_oracle_gemm_blockscaled(a, b, sa, sb, torch.float32)
    cnt = torch.tensor([128.0, 128.0, 44.0])
    want = torch.stack([(sa * sb[0] * cnt).sum(1), (sa * sb[1] * cnt).sum(1)], dim=1)  # [m, 2 column blocks]
    assert torch.equal(c[:, :128], want[:, :1].expand(m, 128)) and torch.equal(c[:, 128:], want[:, 1:].expand(m, 2))
    # random operands: equal to dequantise-then-matmul up to fp32 summation order
    g = torch.Generator().manual_seed(1)
    a = torch.randn(40, 384, generator=g).to(torch.float8_e4m3fn)
    b = torch.randn(140, 384, generator=g).to(torch.float8_e5m2).float().to(torch.float8_e4m3fn)
    sa = torch.rand(40, 3, generator=g) + 0.5
    sb = torch.rand(2, 3, generator=g) + 0.5
    c = R.c_oracle_gemm_blockscaled(a, b, sa, sb, torch.float32)
    torch.testing.assert_close(c, R.blockscaled_fp32(a, b, sa, sb), rtol=1e-4, atol=1e-4)
    cb = R.c_oracle_gemm_blockscaled(a, b, sa, sb, torch.bfloat16)
    assert torch.equal(cb, c.bfloat16())


def test_fused_oracle_order_of_operations():
    """oracle_gemm_fused: ((alpha * acc) * row_scale) * col_scale + bias, each step rounded to fp32, relu, one cast."""
    import numpy as np
    torch.manual_seed(3)
    m, n, k = 9, 11, 24
    a = torch.randint(-3, 4, (m, k)).float().bfloat16()
    b = torch.randint(-3, 4, (n, k)).float().bfloat16()
    rs, cs, bias = torch.rand(m) + 0.5, torch.rand(n) + 0.5, torch.randn(n)
    alpha = 0.37
    got = R.c_oracle_gemm_fused(a, b, torch.float32, alpha=alpha, relu=True, row_scale=rs, col_scale=cs, bias=bias)
    acc = (a.double() @ b.double().t()).numpy().astype(np.float32)  # exact small integers
    v = (acc * np.float32(alpha)).astype(np.float32)
    v = (v * rs.numpy()[:, None]).astype(np.float32)
    v = (v * cs.numpy()[None, :]).astype(np.float32)
    v = (v + bias.numpy()[None, :]).astype(np.float32)
    assert np.array_equal(got.numpy(), np.maximum(v, 0))
    # no vectors = oracle_gemm
    assert torch.equal(R.c_oracle_gemm_fused(a, b, torch.bfloat16, alpha=alpha).view(torch.int16),
                       R.c_oracle_gemm(a, b, torch.bfloat16, alpha=alpha).view(torch.int16))
    per_row = R.c_oracle_gemm_fused(a, b, torch.float32, bias=torch.arange(m).float(), bias_per_row=True)
    assert torch.equal(per_row, torch.from_numpy(acc) + torch.arange(m).float()[:, None])
