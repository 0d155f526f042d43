"""Tensor supply and comparison helpers — same semantics as tilelang/utils/tensor.py."""
from __future__ import annotations

from enum import Enum

import torch


class TensorSupplyType(Enum):  # tilelang/utils/tensor.py:24-31
    Integer = 1
    Uniform = 2
    Normal = 3
    Randn = 4
    Zero = 5
    One = 6
    Auto = 7


def map_torch_type(intype: str) -> torch.dtype:  # tilelang/utils/tensor.py:34-51
    intype = str(intype)
    if intype in ("float8_e4m3", "float8_e4m3fn", "e4m3_float8"):
        return torch.float8_e4m3fn
    if intype in ("float8_e5m2", "e5m2_float8"):
        return torch.float8_e5m2
    if intype == "float":
        return torch.float32
    return getattr(torch, intype)


def is_float8_dtype(dt) -> bool:
    return dt in (torch.float8_e4m3fn, torch.float8_e5m2)


def get_tensor_supply(supply_type: TensorSupplyType = TensorSupplyType.Integer):
    """tilelang/utils/tensor.py:54-133: input generators keyed by supply type."""

    def get_tensor(param) -> torch.Tensor:
        dtype, shape, device = param.dtype, list(map(int, param.shape)), torch.device("cuda", torch.cuda.current_device())
        f8 = is_float8_dtype(dtype)
        if supply_type == TensorSupplyType.Auto:
            if f8:
                return torch.randint(-128, 128, shape, device=device, dtype=torch.int8).to(dtype)
            if dtype in (torch.float16, torch.float32, torch.bfloat16):
                return torch.empty(*shape, device=device, dtype=dtype).uniform_(-1.0, 1.0)
            return torch.randint(-2, 3, shape, device=device, dtype=dtype)
        if supply_type == TensorSupplyType.Integer:
            if f8:
                return torch.randint(-128, 128, shape, device=device, dtype=torch.int8).to(dtype)
            return torch.randint(-2, 3, shape, device=device).to(dtype)
        if supply_type == TensorSupplyType.Uniform:
            return torch.empty(*shape, device=device, dtype=torch.float32).uniform_(-1.0, 1.0).to(dtype)
        if supply_type == TensorSupplyType.Normal:
            return torch.empty(*shape, device=device, dtype=torch.float32).normal_(-1.0, 1.0).to(dtype)
        if supply_type == TensorSupplyType.Randn:
            return torch.randn(*shape, device=device).to(dtype)
        if supply_type == TensorSupplyType.Zero:
            return torch.zeros(*shape, device=device, dtype=dtype)
        if supply_type == TensorSupplyType.One:
            return torch.ones(*shape, device=device, dtype=dtype)
        raise NotImplementedError(supply_type)

    return get_tensor


def torch_assert_close(tensor_a, tensor_b, rtol=1e-2, atol=1e-3, max_mismatched_ratio=0.001, verbose=False,
                       equal_nan=True, check_device=True, check_dtype=True, check_layout=True, check_stride=False,
                       base_name="LHS", ref_name="RHS"):
    """tilelang/utils/tensor.py:217-309: isclose with a bounded ratio of mismatching elements."""
    if check_device and tensor_a.device != tensor_b.device:
        raise AssertionError(f"device mismatch: {tensor_a.device} vs {tensor_b.device}")
    if check_dtype and tensor_a.dtype != tensor_b.dtype:
        raise AssertionError(f"dtype mismatch: {tensor_a.dtype} vs {tensor_b.dtype}")
    if tensor_a.shape != tensor_b.shape:
        raise AssertionError(f"shape mismatch: {tuple(tensor_a.shape)} vs {tuple(tensor_b.shape)}")
    dt = torch.promote_types(tensor_a.dtype, tensor_b.dtype)
    a, b = tensor_a.to(dt), tensor_b.to(dt)
    mismatched = ~torch.isclose(a, b, rtol=rtol, atol=atol, equal_nan=equal_nan)
    num, total = int(mismatched.sum().item()), a.numel()
    allowed = int(total * max_mismatched_ratio)
    if verbose:
        print(f"Number of mismatched elements: {num} / {total} (allowed: {allowed})")
    if num > allowed:
        flat = int(torch.argmax(mismatched.view(-1).int()).item())
        av, bv = a.reshape(-1)[flat].item(), b.reshape(-1)[flat].item()
        diff = (a.float() - b.float()).abs()
        raise AssertionError(
            f"Too many mismatched elements: {num} > {allowed} ({max_mismatched_ratio * 100:.2f}% allowed, but get "
            f"{num / total * 100:.2f}%).\nFirst mismatch at flat index {flat}: {base_name}={av:.6f}, {ref_name}={bv:.6f}"
            f"\nGreatest absolute difference: {diff.max().item()}")
    return True
