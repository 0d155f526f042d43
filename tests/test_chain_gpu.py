"""Two contractions back to back with the intermediate in tensor memory (tlb200_gemm_chain, the "TS" operand mode of
tcgen05.mma: SURVEY §8(f) N4; the reference declares `tl::tcgen05mma_ts`, instruction/tcgen05mma.h:26-133, and leaves
`tcgen5mma_gemm_ts` a TODO, gemm_sm100.h:419-424).  By definition the result equals two ordinary contractions with an
intermediate of the operand type, so the oracle is oracle_gemm applied twice: bit-exact on inputs whose products and
sums are exact, tolerance parity on random data (rel. err <= 1e-2 for 16-bit operands)."""
import ctypes

import pytest
import torch

from oracle import gemm_ref as R
from tilelang_mesh_b200 import _abi
from tilelang_mesh_b200 import gemm as G

pytestmark = pytest.mark.gpu


def _oracle(a, b1, b2, out, alpha1=1.0, alpha2=1.0, relu_mid=False):
    p = R.c_oracle_gemm(a, b1, a.dtype, alpha=alpha1, relu=relu_mid)   # rounded once to the operand type
    return R.c_oracle_gemm(p, b2, out, alpha=alpha2)


@pytest.mark.parametrize("dt", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("m,k,n1,n2", [(128, 64, 64, 128), (300, 136, 128, 200), (256, 72, 256, 384), (77, 320, 192, 40)])
def test_chain_bit_exact_vs_two_oracle_gemms(dt, m, k, n1, n2):
    g = torch.Generator(device="cuda").manual_seed(m + n1)
    a = torch.randint(-2, 3, (m, k), device="cuda", generator=g).float().to(dt)
    b1 = torch.randint(-2, 3, (n1, k), device="cuda", generator=g).float().to(dt)
    b2 = torch.randint(-2, 3, (n2, n1), device="cuda", generator=g).float().to(dt)
    # |S| <= 4 * k: with alpha1 = 2^-3 the intermediate is a multiple of 1/8 below 256 -> exact in bf16 and fp16
    for out in (torch.float32, dt):
        got = G.gemm_chain(a, b1, b2, out_dtype=out, alpha1=0.125, alpha2=0.5)
        want = _oracle(a, b1, b2, out, 0.125, 0.5)
        assert torch.equal(got.cpu().float(), want.float()), (m, k, n1, n2, out)
    got = G.gemm_chain(a, b1, b2, out_dtype=torch.float32, alpha1=0.125, relu_mid=True)
    assert torch.equal(got.cpu(), _oracle(a, b1, b2, torch.float32, 0.125, 1.0, True))


def test_chain_rounds_the_intermediate_once_to_the_operand_type():
    # values that are NOT representable in bf16 after the first contraction: the oracle rounds them (RNE), so must we
    m, k, n1, n2 = 128, 64, 64, 128
    a = torch.full((m, k), 1.0, device="cuda").bfloat16()
    b1 = (torch.arange(n1, device="cuda").float()[:, None] * 0 + 1.0).expand(n1, k).contiguous().bfloat16()
    b1[:, 0] = 1.0078125  # 1 + 2^-7: sum over k = 64.0078125, not a bf16 value
    b2 = torch.eye(n1, device="cuda").bfloat16()[:n2 if n2 < n1 else n1].contiguous()
    b2 = torch.cat([b2, torch.zeros(n2 - b2.shape[0], n1, device="cuda").bfloat16()])
    got = G.gemm_chain(a, b1, b2, out_dtype=torch.float32)
    want = _oracle(a, b1, b2, torch.float32)
    assert torch.equal(got.cpu(), want)
    assert got[0, 0].item() == 64.0  # 64.0078125 -> bf16 64.0


def test_chain_tolerance_on_random_data_and_many_row_tiles():
    m, k, n1, n2 = 20000, 1024, 128, 1000   # a low-rank product x A^T B^T; more row tiles than SMs, ragged N2
    a = torch.randn(m, k, device="cuda").bfloat16()
    b1 = (torch.randn(n1, k, device="cuda") / 32).bfloat16()
    b2 = (torch.randn(n2, n1, device="cuda") / 8).bfloat16()
    got = G.gemm_chain(a, b1, b2)
    p = (a.float() @ b1.float().t()).bfloat16()
    want = (p.float() @ b2.float().t()).bfloat16()
    R.assert_allclose(got, want, rtol=1e-2, atol=1e-2)
    torch.cuda.synchronize()
    # run-to-run bit equality
    assert torch.equal(got, G.gemm_chain(a, b1, b2))


def test_chain_argument_errors():
    lib = _abi.load()
    a = torch.zeros(128, 64, device="cuda", dtype=torch.bfloat16)
    with pytest.raises(RuntimeError, match="64, 128, 192 or 256"):
        G.gemm_chain(a, torch.zeros(96, 64, device="cuda", dtype=torch.bfloat16), torch.zeros(128, 96, device="cuda", dtype=torch.bfloat16))
    with pytest.raises(ValueError, match="K shape check"):
        G.gemm_chain(a, torch.zeros(64, 32, device="cuda", dtype=torch.bfloat16), torch.zeros(128, 64, device="cuda", dtype=torch.bfloat16))
    with pytest.raises(ValueError, match="float16 / bfloat16"):
        G.gemm_chain(a.float(), torch.zeros(64, 64, device="cuda"), torch.zeros(128, 64, device="cuda"))
    d = _abi.ChainDesc()
    d.abi_version = 1
    assert lib.tlb200_gemm_chain(ctypes.byref(d), None, None, None, None, None) != 0 and "abi_version" in _abi.last_error()
