"""`tilelang.jit` / `tilelang.compile` / `JITKernel` — the reference's JIT surface
(tilelang/jit/__init__.py:48-119,456-540, tilelang/jit/kernel.py:31-611) bound to the C-ABI shim.

"Compilation" here is: trace the tile program, recognise the GEMM family (program.py), resolve a
tile plan in libtlb200.so.  The arithmetic runs in the ahead-of-time built sm_100a kernel; nothing
is generated or compiled per program, and nothing falls back to PyTorch or the CPU.
"""
from __future__ import annotations

import inspect
import logging
from typing import Any, Callable, Dict, List, Optional

import torch

from .. import language as T
from ..language.expr import Expr, Var
from ..program import BlockScaledProgram, GemmProgram, recognize
from ..utils.target import determine_target
from ..utils.tensor import TensorSupplyType, map_torch_type

logger = logging.getLogger("tilelang_mesh_b200")

_BACKENDS = ("auto", "dlpack", "tvm_ffi", "ctypes", "cython", "nvrtc", "torch")


class KernelParam:
    """tilelang/engine/param.py: dtype + static shape of one kernel parameter."""

    def __init__(self, name: str, dtype: torch.dtype, shape):
        # dynamic dimensions (T.dynamic / tvm.te.var) stay symbolic and are bound from the tensors of each call
        self.name, self.dtype, self.shape = name, dtype, tuple(s if isinstance(s, Expr) else int(s) for s in shape)

    @property
    def is_dynamic(self) -> bool:
        return any(isinstance(s, Expr) for s in self.shape)

    def __repr__(self):
        return f"KernelParam({self.name}: {self.dtype}{list(self.shape)})"


def _legalize_out_idx(out_idx, n_params: int) -> List[int]:
    """tilelang/jit/adapter/base.py:20-44: int or list, negatives count from the end."""
    if out_idx is None:
        return []
    idx = [out_idx] if isinstance(out_idx, int) else list(out_idx)
    res = []
    for i in idx:
        if not isinstance(i, int):
            raise ValueError("result_idx should be a list of integers")
        if i >= n_params or i < -n_params:
            raise ValueError(f"The output index should be less than the number of parameters, but got {i}")
        res.append(i + n_params if i < 0 else i)
    return res


class JITKernel:
    """A compiled tile program: callable on torch tensors (tilelang/jit/kernel.py:31)."""

    def __init__(self, func: "T.PrimFunc", out_idx=None, execution_backend: str = "auto", target="auto",
                 target_host=None, verbose: bool = False, pass_configs: Optional[Dict[str, Any]] = None,
                 compile_flags=None, tile_hints: Optional[dict] = None):
        if not isinstance(func, T.PrimFunc):
            raise TypeError(f"tilelang.compile expects a @T.prim_func, got {type(func).__name__}")
        if execution_backend not in _BACKENDS:
            raise ValueError(f"Invalid execution backend. {execution_backend!r} not in {_BACKENDS}")
        self.prim_func = func
        self.target = determine_target(target)
        self.execution_backend = "ctypes"  # one adapter: ctypes over the C ABI
        self.verbose, self.pass_configs, self.compile_flags = verbose, dict(pass_configs or {}), compile_flags
        self._params = [KernelParam(n, map_torch_type(a.dtype), a.shape) for n, a in func.params]
        self._dynamic = any(kp.is_dynamic for kp in self._params)
        self._programs: Dict[tuple, Any] = {}
        if self._dynamic:
            # Validate the program's structure once with every shape symbol at a nominal size; the plan itself is
            # resolved per call from the actual tensor shapes (examples/dynamic_shape/example_dynamic.py).
            syms = []
            for kp in self._params:
                for d in kp.shape:
                    if isinstance(d, Expr) and not isinstance(d, Var):
                        raise NotImplementedError("a dynamic dimension must be a plain shape symbol (T.dynamic / te.var)")
                    if isinstance(d, Var) and not any(d is x for x in syms):
                        syms.append(d)
            self._symbols = syms
            self.program = recognize(func, {v: 1024 for v in syms})
        else:
            self.program = recognize(func)  # GemmProgram or BlockScaledProgram
        self._out_idx = _legalize_out_idx(out_idx, len(self._params))
        if any(i != self.program.c_param for i in self._out_idx):
            raise ValueError(f"out_idx {self._out_idx} names an input tensor; the program's only output is parameter "
                             f"{self.program.c_param}")
        self.tile_hints = dict(tile_hints or {})
        self.latency: Optional[float] = None
        self.config: Optional[dict] = None
        self.ref_latency: Optional[float] = None
        for n in self.program.notes:
            logger.info(n)
        if verbose:
            logger.info("compiled %s", self.program.summary())

    @classmethod
    def from_tilelang_function(cls, tilelang_func, **kwargs):
        return cls(tilelang_func, **kwargs)

    # ---------------------------------------------------------------- call
    @property
    def needs_outputs(self) -> bool:
        return not self._out_idx

    def __call__(self, *args):
        """tilelang/jit/adapter/cython/cython_wrapper.pyx:164-287 semantics: inputs only when out_idx is
        set (outputs are allocated with torch.empty and returned), otherwise every parameter is passed."""
        from .. import gemm as G
        p = self.program
        n_in = len(self._params) - len(self._out_idx)
        if len(args) != n_in:
            raise ValueError(f"Expected {len(self._params)} inputs, got {len(args) + len(self._out_idx)} with "
                             f"{len(args)} inputs and {len(self._out_idx)} outputs")
        full: List[Optional[torch.Tensor]] = []
        it = iter(args)
        for i in range(len(self._params)):
            full.append(None if i in self._out_idx else next(it))
        env: Dict[Var, int] = {}
        for i, t in enumerate(full):
            if t is None:
                continue
            kp = self._params[i]
            if not isinstance(t, torch.Tensor):
                raise ValueError(f"argument {i} ({kp.name}) must be a torch.Tensor")
            if t.dtype != kp.dtype:
                raise ValueError(f"argument {i} ({kp.name}) dtype mismatch: expected {kp.dtype}, got {t.dtype}")
            ok = t.dim() == len(kp.shape)
            for want, got in zip(kp.shape, t.shape) if ok else ():
                if isinstance(want, Var):  # bind the symbol on first sight, then insist on the same value
                    ok = ok and env.setdefault(want, int(got)) == int(got)
                else:
                    ok = ok and want == got
            if not ok:
                raise ValueError(f"argument {i} ({kp.name}) shape mismatch: expected {list(kp.shape)}, got {list(t.shape)}")
        if self._dynamic:
            missing = [v.name for v in self._symbols if not any(v is k for k in env)]
            if missing:
                raise ValueError(f"shape symbols {missing} appear only in output tensors: pass those tensors explicitly")
            key = tuple(env[v] for v in self._symbols)
            p = self._programs.get(key)
            if p is None:
                p = self._programs[key] = recognize(self.prim_func, env)
        a, b, c = full[p.a_param], full[p.b_param], full[p.c_param]
        if c is None:
            shape = tuple(env[d] if isinstance(d, Var) else d for d in self._params[p.c_param].shape)
            c = torch.empty(shape, dtype=self._params[p.c_param].dtype, device=a.device)
        if isinstance(p, BlockScaledProgram):
            hints = {k: v for k, v in self.tile_hints.items() if k in ("cta_group", "block_n", "stages", "raster")}
            G.gemm_blockscaled(a, b, full[p.sa_param], full[p.sb_param], c, **hints)
        else:
            extra = dict(self.tile_hints, accumulate=True) if p.accumulate else self.tile_hints
            G.gemm(a, b, c, trans_a=p.trans_A, trans_b=p.trans_B, alpha=p.alpha, relu=p.relu, **extra)
        return c if self._out_idx else None

    # ---------------------------------------------------------------- introspection
    @property
    def out_idx(self) -> List[int]:
        return self._out_idx

    @property
    def params(self) -> List[KernelParam]:
        return self._params

    def plan(self) -> dict:
        from .. import gemm as G
        p = self.program
        if isinstance(p, BlockScaledProgram):  # fixed plan: tlb200_gemm_blockscaled's default (tlb200.cu)
            sms = torch.cuda.get_device_properties(0).multi_processor_count if torch.cuda.is_available() else 148
            bn = self.tile_hints.get("block_n") or (128 if p.N <= 128 or -(-p.M // 128) * -(-p.N // 256) < sms else 256)
            cg = self.tile_hints.get("cta_group") or (2 if bn == 256 and p.M > 128 and -(-p.M // 256) * -(-p.N // bn) >= sms // 2 else 1)
            return dict(kernel="gemm_blockscaled_kernel", cta_group=cg, block_m=128 * cg, block_n=bn, block_k=128,
                        tiles_m=-(-p.M // (128 * cg)), tiles_n=-(-p.N // bn), grid=1, threads=384, tmem_cols=2 * bn)
        return G.plan(p.M, p.N, p.K, map_torch_type(p.in_dtype), map_torch_type(p.out_dtype), p.trans_A, p.trans_B,
                      **self.tile_hints)

    def get_kernel_source(self, kernel_only: bool = True) -> str:
        """The reference returns generated CUDA; this backend is ahead-of-time, so it returns the
        resolved plan followed by the hand-written kernel's source."""
        import os
        from .. import build
        hdr = "// " + self.program.summary() + "\n// plan: " + repr(self.plan()) + "\n"
        main = "gemm_blockscaled_sm100.cuh" if isinstance(self.program, BlockScaledProgram) else "gemm_sm100.cuh"
        srcs = [os.path.join(build.CSRC, f) for f in (main,) + (() if kernel_only else ("ptx.cuh", "tlb200.cu"))]
        return hdr + "\n".join(open(s).read() for s in srcs)

    @property
    def kernel_source(self) -> str:
        return self.get_kernel_source()

    def get_host_source(self) -> str:
        import os
        from .. import build
        return open(os.path.join(build.CSRC, "tlb200.cu")).read()

    def get_profiler(self, tensor_supply_type: TensorSupplyType = TensorSupplyType.Auto):
        from ..profiler import Profiler
        return Profiler(self._params, self._out_idx, tensor_supply_type).with_default_adapter(self)

    def run_once(self, func: Optional[Callable] = None):
        return self.get_profiler().run_once(func)

    def update_tuner_result(self, latency: float, config: dict, ref_latency: Optional[float]):
        self.latency, self.config, self.ref_latency = latency, config, ref_latency
        return self

    def get_tuner_result(self) -> dict:
        return {"latency": self.latency, "config": self.config, "ref_latency": self.ref_latency}


def compile(func=None, out_idx=None, execution_backend: str = "auto", target="auto", target_host=None,  # noqa: A001
            verbose: bool = False, pass_configs: Optional[Dict[str, Any]] = None, compile_flags=None) -> JITKernel:
    """tilelang.compile — tilelang/jit/__init__.py:48-119."""
    return JITKernel(func, out_idx=out_idx, execution_backend=execution_backend, target=target, target_host=target_host,
                     verbose=verbose, pass_configs=pass_configs, compile_flags=compile_flags)


def par_compile(funcs, out_idx=None, **kw) -> List[JITKernel]:
    return [compile(f, out_idx=out_idx, **kw) for f in funcs]


class JITImpl:
    """What `@tilelang.jit` returns: calling it with the factory's arguments yields a cached JITKernel
    (tilelang/jit/__init__.py:260-430)."""

    def __init__(self, func, out_idx, execution_backend, target, target_host, verbose, pass_configs, debug_root_path,
                 compile_flags):
        self.func, self.out_idx, self.execution_backend = func, out_idx, execution_backend
        self.target, self.target_host, self.verbose = target, target_host, verbose
        self.pass_configs, self.debug_root_path, self.compile_flags = pass_configs, debug_root_path, compile_flags
        self._kernel_cache: Dict[tuple, JITKernel] = {}
        orig = func.orig_func if isinstance(func, T.PrimFunc) else func
        self.signature = inspect.signature(orig)
        self.__name__ = getattr(orig, "__name__", "kernel")
        self.__doc__ = getattr(orig, "__doc__", None)

    def get_tir(self, *args, **kwargs) -> "T.PrimFunc":
        if isinstance(self.func, T.PrimFunc):
            return self.func
        tir = self.func(*args, **kwargs)
        if not isinstance(tir, T.PrimFunc):
            raise ValueError(f"Invalid function type: {type(tir)}; a @tilelang.jit factory must return a @T.prim_func")
        return tir

    def compile(self, *args, **kwargs) -> JITKernel:
        return compile(self.get_tir(*args, **kwargs), out_idx=self.out_idx, execution_backend=self.execution_backend,
                       target=self.target, target_host=self.target_host, verbose=self.verbose,
                       pass_configs=self.pass_configs, compile_flags=self.compile_flags)

    def __call__(self, *args, **kwargs):
        tune_params = kwargs.pop("__tune_params", {})
        key = (args, tuple(sorted(kwargs.items())), tuple(sorted(tune_params.items())))
        try:
            kernel = self._kernel_cache.get(key)
        except TypeError:  # unhashable argument: compile without caching
            return self.compile(*args, **kwargs, **tune_params)
        if kernel is None:
            kernel = self.compile(*args, **kwargs, **tune_params)
            self._kernel_cache[key] = kernel
        return kernel


def jit(func=None, *, out_idx=None, target="auto", target_host=None, execution_backend="auto", verbose=False,
        pass_configs=None, debug_root_path=None, compile_flags=None):
    """tilelang.jit — usable bare (`@tilelang.jit`) or configured (`@tilelang.jit(out_idx=[2])`),
    tilelang/jit/__init__.py:456-540."""
    if isinstance(compile_flags, str):
        compile_flags = [compile_flags]

    def decorator(f):
        return JITImpl(f, out_idx, execution_backend, target, target_host, verbose, pass_configs, debug_root_path,
                       compile_flags)

    if func is not None:
        if not (callable(func) or isinstance(func, T.PrimFunc)):
            raise TypeError("@tilelang.jit must decorate a function or a prim_func")
        return decorator(func)
    return decorator
