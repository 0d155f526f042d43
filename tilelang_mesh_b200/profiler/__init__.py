"""Profiler (tilelang/profiler/__init__.py, bench.py): same timing discipline as the reference's
do_bench — L2 flushed with a 256 MB write before every timed call, per-call CUDA events, mean."""
from __future__ import annotations

from functools import partial
from typing import Callable, List, Optional

import torch

from ..utils.tensor import TensorSupplyType, get_tensor_supply, is_float8_dtype, torch_assert_close


def do_bench(fn: Callable, warmup: float = 25, rep: float = 100, _n_warmup: int = 0, _n_repeat: int = 0,
             quantiles: Optional[List[float]] = None, fast_flush: bool = True, backend: str = "event",
             return_mode: str = "mean"):
    """tilelang/profiler/bench.py:63-170 (CUDA-event backend)."""
    assert return_mode in ("min", "max", "mean", "median"), f"Invalid return_mode: {return_mode}"
    fn()
    torch.cuda.synchronize()
    cache = torch.empty(int(256e6 // 4), dtype=torch.int, device="cuda") if fast_flush else torch.empty(
        int(256e6), dtype=torch.int8, device="cuda")
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(5):
        cache.zero_()
        fn()
    e.record()
    torch.cuda.synchronize()
    estimate_ms = s.elapsed_time(e) / 5
    n_warmup = _n_warmup if _n_warmup > 0 else max(1, int(warmup / estimate_ms))
    n_repeat = _n_repeat if _n_repeat > 0 else max(1, int(rep / estimate_ms))
    for _ in range(n_warmup):
        fn()
    starts = [torch.cuda.Event(enable_timing=True) for _ in range(n_repeat)]
    ends = [torch.cuda.Event(enable_timing=True) for _ in range(n_repeat)]
    for i in range(n_repeat):
        cache.zero_()
        starts[i].record()
        fn()
        ends[i].record()
    torch.cuda.synchronize()
    times = torch.tensor([a.elapsed_time(b) for a, b in zip(starts, ends)], dtype=torch.float)
    if quantiles is not None:
        r = torch.quantile(times, torch.tensor(quantiles, dtype=torch.float)).tolist()
        return r[0] if len(r) == 1 else r
    return getattr(torch, return_mode)(times).item()


class Profiler:
    """tilelang/profiler/__init__.py:21-280, for kernels of this backend."""

    def __init__(self, params, result_idx, supply_type: TensorSupplyType = TensorSupplyType.Auto, adapter=None):
        self.params, self.result_idx, self.adapter = params, list(result_idx or []), adapter
        self.supply_type = supply_type
        self.supply = get_tensor_supply(supply_type)

    def with_default_adapter(self, adapter):
        self.adapter = adapter
        return self

    def _get_inputs(self, with_output: bool = False):
        return [self.supply(p) for i, p in enumerate(self.params) if with_output or i not in self.result_idx]

    @property
    def func(self):
        assert self.adapter is not None, "adapter should be provided"
        return self.adapter

    def __call__(self, *args, **kwds):
        return self.func(*args, **kwds)

    def run_once(self, func: Optional[Callable] = None):
        ins = self._get_inputs()
        return (func or self.func)(*ins)

    def assert_allclose(self, reference_program: Callable, input_tensors=None, atol: float = 1e-2, rtol: float = 1e-2,
                        max_mismatched_ratio=0.01):
        ins = self._get_inputs() if input_tensors is None else input_tensors
        ref_outs = reference_program(*ins)
        torch.cuda.synchronize()
        lib_outs = self.func(*ins)
        torch.cuda.synchronize()
        norm = lambda o: [o] if isinstance(o, torch.Tensor) else list(o) if isinstance(o, (tuple, list)) else []  # noqa: E731
        ref_outs, lib_outs = norm(ref_outs), norm(lib_outs)
        assert len(lib_outs) == len(ref_outs), "len(lib_tensors) not equals to len(ref_tensors) !"
        for lhs, rhs in zip(lib_outs, ref_outs):
            if lhs is not None and rhs is not None:
                torch_assert_close(lhs.float() if is_float8_dtype(lhs.dtype) else lhs,
                                   rhs.float() if is_float8_dtype(rhs.dtype) else rhs, rtol=rtol, atol=atol,
                                   max_mismatched_ratio=max_mismatched_ratio, base_name="tilelang", ref_name="ref",
                                   check_dtype=False)

    def assert_consistent(self, repeat: int = 10):
        ins = self._get_inputs()
        ref = self.func(*ins)
        for _ in range(repeat):
            out = self.func(*ins)
            for a, b in zip([ref] if isinstance(ref, torch.Tensor) else ref, [out] if isinstance(out, torch.Tensor) else out):
                assert torch.equal(a.view(torch.uint8), b.view(torch.uint8)), "result is not consistent across runs"

    def do_bench(self, func: Optional[Callable] = None, warmup: int = 25, rep: int = 100, n_warmup: int = 1,
                 n_repeat: int = 1, input_tensors=None, backend: str = "event", quantiles=None,
                 return_mode: str = "mean") -> float:
        if func is None:
            func = self.func
            ins = self._get_inputs(with_output=getattr(func, "needs_outputs", False)) if input_tensors is None else input_tensors
        else:
            ins = self._get_inputs() if input_tensors is None else input_tensors
        return do_bench(partial(func, *ins), warmup=warmup, rep=rep, _n_warmup=n_warmup, _n_repeat=n_repeat,
                        quantiles=quantiles, backend=backend, return_mode=return_mode)
