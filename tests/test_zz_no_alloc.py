"""The allocation-free entry point, tlb200_gemm_static_scratch.

Some users of this library cannot get device memory for it: a wrapper that has to keep the
reference's generated `call(A, B, C, stream)` signature is given tensors and a stream only, and
the framework above it owns all device memory through its own pool.  For them split-K's partial
sums live in a fixed scratch array inside the library instead of a caller-supplied workspace.

The test drives the entry point through the C ABI, twice in a row on different data (the scratch
keeps the first call's contents, which must not leak into the second), and compares each result
with what the workspace form (tlb200_gemm_ex) returns for the same inputs — same tolerance as the
other GPU tests, rel. err <= 1e-2 for bf16.  It lives in its own file, collected last, and must
pass both alone in a fresh process and after every other test has run.
"""
import ctypes

import pytest
import torch

from oracle import gemm_ref as R
from tilelang_mesh_b200 import _abi
from tilelang_mesh_b200 import gemm as G

pytestmark = pytest.mark.gpu

M, N, K = 128, 4096, 14336  # decode-like: 16-32 output tiles for 148 SMs -> the plan splits K


def _inputs(seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    a = torch.randn(M, K, device="cuda", generator=g).bfloat16()
    b = torch.randn(N, K, device="cuda", generator=g).bfloat16()
    return a, b


def test_static_scratch_matches_workspace_form_twice_in_a_row():
    lib = _abi.load()
    G._ensure_init(torch.cuda.current_device())
    stream = torch.cuda.current_stream().cuda_stream
    desc = G.make_desc(M, N, K, _abi.BFLOAT16, _abi.BFLOAT16, 0, 1, K, K, N)
    need = lib.tlb200_gemm_workspace_bytes(ctypes.byref(desc))
    assert 0 < need <= _abi.STATIC_SCRATCH_BYTES, "the chosen shape must exercise split-K inside the static limit"
    ws = torch.empty(need, dtype=torch.uint8, device="cuda")
    data = [_inputs(1), _inputs(2)]
    got = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    want = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
    for call, (a, b) in enumerate(data):
        got.fill_(float("nan"))
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info()[0]
        rc = lib.tlb200_gemm_static_scratch(ctypes.byref(desc), a.data_ptr(), b.data_ptr(), got.data_ptr(), stream)
        assert rc == 0, _abi.last_error()
        torch.cuda.synchronize()
        if call == 1:  # (the very first launch of a process may still load the kernel image)
            assert torch.cuda.mem_get_info()[0] == free_before, "the call changed the amount of free device memory"
        rc = lib.tlb200_gemm_ex(ctypes.byref(desc), a.data_ptr(), b.data_ptr(), want.data_ptr(), ws.data_ptr(), need, stream)
        assert rc == 0, _abi.last_error()
        torch.cuda.synchronize()
        R.assert_allclose(got, want, rtol=1e-2, atol=1e-2)          # the two entry points agree
        R.assert_allclose(got, R.matmul_fp32(a, b).bfloat16(), rtol=1e-2, atol=1e-2)  # and both match the oracle
        assert not torch.isnan(got.float()).any()


def test_static_scratch_rejects_plans_beyond_its_size():
    lib = _abi.load()
    G._ensure_init(torch.cuda.current_device())
    m, n, k = 256, 8192, 16384
    desc = G.make_desc(m, n, k, _abi.BFLOAT16, _abi.BFLOAT16, 0, 1, k, k, n, cta_group=2, block_n=256, split_k=8)
    assert lib.tlb200_gemm_workspace_bytes(ctypes.byref(desc)) > _abi.STATIC_SCRATCH_BYTES
    a = torch.zeros(m, k, device="cuda", dtype=torch.bfloat16)
    b = torch.zeros(n, k, device="cuda", dtype=torch.bfloat16)
    c = torch.full((m, n), 3.0, device="cuda", dtype=torch.bfloat16)
    rc = lib.tlb200_gemm_static_scratch(ctypes.byref(desc), a.data_ptr(), b.data_ptr(), c.data_ptr(), None)
    assert rc != 0
    msg = _abi.last_error()
    assert "static" in msg and str(_abi.STATIC_SCRATCH_BYTES) in msg
    torch.cuda.synchronize()
    assert (c == 3).all(), "a rejected call must not touch the output"


def test_static_scratch_without_split_behaves_like_plain_gemm():
    lib = _abi.load()
    G._ensure_init(torch.cuda.current_device())
    m = n = k = 1024
    desc = G.make_desc(m, n, k, _abi.BFLOAT16, _abi.BFLOAT16, 0, 1, k, k, n)
    a = torch.randint(-3, 4, (m, k), device="cuda").float().bfloat16()
    b = torch.randint(-3, 4, (n, k), device="cuda").float().bfloat16()
    c = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
    assert lib.tlb200_gemm_static_scratch(ctypes.byref(desc), a.data_ptr(), b.data_ptr(), c.data_ptr(),
                                          torch.cuda.current_stream().cuda_stream) == 0, _abi.last_error()
    assert torch.equal(c, G.gemm(a, b))
