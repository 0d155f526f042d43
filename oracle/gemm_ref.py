"""Numerical oracles for the contraction path (TEST INFRASTRUCTURE ONLY — see oracle/__init__.py).

 * matmul_fp32      — the primary oracle north_star names: fp32 torch.matmul of the exact
                      low-precision inputs up-cast to fp32 (reference tests do the same at run
                      time: testing/python/kernel/test_tilelang_kernel_gemm.py:88-102,
                      examples/gemm_sm100/gemm_tcgen5mma.py:82-86).
 * c_oracle_gemm    — ctypes call into oracle/gemm_oracle.c (scalar C restatement).
 * assert_allclose  — the reference's comparison rule: torch.isclose(rtol, atol) with a bounded
                      ratio of mismatching elements (tilelang/utils/tensor.py:217-309;
                      Profiler.assert_allclose defaults atol=rtol=1e-2, max_mismatched_ratio=0.01,
                      tilelang/profiler/__init__.py:76-83).
 * calc_diff        — 1 - 2<x,y>/(|x|^2+|y|^2), the fp8 tests' metric
                      (testing/python/kernel/test_tilelang_kernel_fp8_gemm.py:8-12, threshold 1e-3).
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np
import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "liboracle_gemm.so")

DT = {torch.float16: 0, torch.bfloat16: 1, torch.float8_e4m3fn: 2, torch.float8_e5m2: 3, torch.float32: 4}

# north_star tolerances
RTOL = {torch.float16: 1e-2, torch.bfloat16: 1e-2, torch.float8_e4m3fn: 5e-2, torch.float8_e5m2: 5e-2}


def build_c_oracle(force: bool = False) -> str:
    src = os.path.join(_HERE, "gemm_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.run(["make", "-C", _HERE, "-s", "-B" if force else "-s"], check=True)
    return _SO


_lib = None


def _c():
    global _lib
    if _lib is None:
        lib = ctypes.CDLL(build_c_oracle())
        lib.oracle_gemm.restype = ctypes.c_int
        lib.oracle_gemm.argtypes = [ctypes.c_int] * 7 + [ctypes.c_void_p, ctypes.c_longlong, ctypes.c_void_p,
                                                         ctypes.c_longlong, ctypes.c_void_p, ctypes.c_longlong,
                                                         ctypes.c_float, ctypes.c_int, ctypes.c_int]
        lib.oracle_decode.restype = ctypes.c_float
        lib.oracle_decode.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_longlong]
        lib.oracle_encode.restype = None
        lib.oracle_encode.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_longlong, ctypes.c_float]
        _lib = lib
    return _lib


def c_oracle_gemm(a: torch.Tensor, b: torch.Tensor, out_dtype: torch.dtype, trans_a=False, trans_b=True,
                  alpha: float = 1.0, relu: bool = False, product_mode: int = 0) -> torch.Tensor:
    """Scalar C oracle on CPU tensors (small sizes only: O(MNK) scalar loop)."""
    a, b = a.detach().cpu().contiguous(), b.detach().cpu().contiguous()
    m, k = (a.shape[1], a.shape[0]) if trans_a else a.shape
    n = b.shape[0] if trans_b else b.shape[1]
    c = torch.empty((m, n), dtype=out_dtype)
    rc = _c().oracle_gemm(m, n, k, DT[a.dtype], DT[out_dtype], int(trans_a), int(trans_b), a.data_ptr(),
                          a.stride(0) if a.shape[0] > 1 else a.shape[1], b.data_ptr(),
                          b.stride(0) if b.shape[0] > 1 else b.shape[1], c.data_ptr(), n, float(alpha), int(relu),
                          int(product_mode))
    if rc != 0:
        raise ValueError("oracle_gemm: bad arguments")
    return c


def c_oracle_gemm_fused(a: torch.Tensor, b: torch.Tensor, out_dtype: torch.dtype, trans_a=False, trans_b=True, alpha: float = 1.0,
                        relu: bool = False, row_scale=None, col_scale=None, bias=None, bias_per_row: bool = False) -> torch.Tensor:
    """Scalar C oracle of the contraction with epilogue vectors (oracle_gemm_fused; small sizes only)."""
    a, b = a.detach().cpu().contiguous(), b.detach().cpu().contiguous()
    m, k = (a.shape[1], a.shape[0]) if trans_a else a.shape
    n = b.shape[0] if trans_b else b.shape[1]
    c = torch.empty((m, n), dtype=out_dtype)
    vecs = [None if v is None else v.detach().cpu().float().contiguous() for v in (row_scale, col_scale, bias)]
    fn = _c().oracle_gemm_fused
    fn.restype = ctypes.c_int
    fn.argtypes = [ctypes.c_int] * 7 + [ctypes.c_void_p, ctypes.c_longlong, ctypes.c_void_p, ctypes.c_longlong, ctypes.c_void_p,
                                        ctypes.c_longlong, ctypes.c_float, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p,
                                        ctypes.c_void_p, ctypes.c_int]
    rc = fn(m, n, k, DT[a.dtype], DT[out_dtype], int(trans_a), int(trans_b), a.data_ptr(),
            a.stride(0) if a.shape[0] > 1 else a.shape[1], b.data_ptr(), b.stride(0) if b.shape[0] > 1 else b.shape[1],
            c.data_ptr(), n, float(alpha), int(relu), *[None if v is None else v.data_ptr() for v in vecs], int(bias_per_row))
    if rc != 0:
        raise ValueError("oracle_gemm_fused: bad arguments")
    return c


def c_oracle_gemm_blockscaled(a: torch.Tensor, b: torch.Tensor, scales_a: torch.Tensor, scales_b: torch.Tensor,
                              out_dtype: torch.dtype) -> torch.Tensor:
    """Scalar C oracle of the fine-grained-scale fp8 contraction (small sizes only)."""
    a, b = a.detach().cpu().contiguous(), b.detach().cpu().contiguous()
    sa, sb = scales_a.detach().cpu().float().contiguous(), scales_b.detach().cpu().float().contiguous()
    m, k = a.shape
    n = b.shape[0]
    c = torch.empty((m, n), dtype=out_dtype)
    fn = _c().oracle_gemm_blockscaled
    fn.argtypes = [ctypes.c_int] * 5 + [ctypes.c_void_p, ctypes.c_longlong, ctypes.c_void_p, ctypes.c_longlong,
                                        ctypes.c_void_p, ctypes.c_longlong, ctypes.c_void_p, ctypes.c_longlong,
                                        ctypes.c_longlong, ctypes.c_void_p, ctypes.c_longlong, ctypes.c_longlong]
    fn.restype = ctypes.c_int
    rc = fn(m, n, k, DT[a.dtype], DT[out_dtype], a.data_ptr(), max(k, 1), b.data_ptr(), max(k, 1), c.data_ptr(), n,
            sa.data_ptr(), sa.stride(0) if sa.numel() else 0, 1, sb.data_ptr(), sb.stride(0) if sb.numel() else 0, 1)
    if rc != 0:
        raise ValueError("oracle_gemm_blockscaled: bad arguments")
    return c


def blockscaled_fp32(a: torch.Tensor, b: torch.Tensor, scales_a: torch.Tensor, scales_b: torch.Tensor) -> torch.Tensor:
    """The same contraction in fp32 torch ops (any size): dequantise, then one fp32 matmul.  Equal to the
    oracle up to fp32 summation order; this is what the reference's checker computes per 128x128 block with
    torch._scaled_mm (example_deepgemm_fp8_2xAcc.py:109-135), without its bf16 rounding of every partial."""
    m, k = a.shape
    n = b.shape[0]
    af = a.float() * scales_a.float().repeat_interleave(128, dim=1)[:, :k]
    bf = b.float() * scales_b.float().repeat_interleave(128, dim=0)[:n].repeat_interleave(128, dim=1)[:, :k]
    return af @ bf.t()


def matmul_fp32(a: torch.Tensor, b: torch.Tensor, trans_a=False, trans_b=True, alpha: float = 1.0,
                relu: bool = False) -> torch.Tensor:
    """fp32 reference result (not yet cast to the output type)."""
    af = a.float().t() if trans_a else a.float()
    bf = b.float().t() if trans_b else b.float()
    c = af @ bf
    if alpha != 1.0:
        c = c * alpha
    if relu:
        c = torch.relu(c)
    return c


def assert_allclose(actual: torch.Tensor, expected: torch.Tensor, rtol=1e-2, atol=1e-2,
                    max_mismatched_ratio=0.01, name="result"):
    a, e = actual.float(), expected.float()
    if a.shape != e.shape:
        raise AssertionError(f"{name}: shape {tuple(a.shape)} != {tuple(e.shape)}")
    mism = ~torch.isclose(a, e, rtol=rtol, atol=atol, equal_nan=True)
    bad, total = int(mism.sum().item()), a.numel()
    if bad > int(total * max_mismatched_ratio):
        idx = mism.nonzero()[0].tolist()
        raise AssertionError(
            f"{name}: too many mismatched elements: {bad} > {int(total * max_mismatched_ratio)} of {total}; first at "
            f"{idx}: got {a[tuple(idx)].item()} expected {e[tuple(idx)].item()}; greatest abs diff "
            f"{(a - e).abs().max().item()}")
    return bad


def max_rel_err(actual: torch.Tensor, expected_fp32: torch.Tensor) -> float:
    """Greatest |got - ref| relative to the tensor's scale (Frobenius-normalised per element).

    north_star's 'rel. err <= 1e-2 / 5e-2 vs an fp32 torch.matmul oracle': element errors are
    measured against max(|ref|, rms(ref)) so near-zero outputs of a random GEMM do not blow up
    the ratio."""
    a, e = actual.double(), expected_fp32.double()
    scale = torch.maximum(e.abs(), torch.sqrt((e * e).mean()).clamp_min(1e-30))
    return float(((a - e).abs() / scale).max().item())


def calc_diff(x: torch.Tensor, y: torch.Tensor) -> float:
    x, y = x.double(), y.double()
    den = (x * x + y * y).sum()
    if den == 0:
        return 0.0
    return float(1 - 2 * (x * y).sum() / den)
