"""`T` — the tile-language surface of the reference, as a tracing front-end.

The reference's `tilelang.language` builds TVM TIR; its GEMM programs reach the hot path through
two tile operators, `T.gemm` (tilelang/language/gemm.py:105-163) and `T.copy`
(tilelang/language/copy.py:13-94).  This module keeps those operator signatures (names, argument
order, defaults) and the handful of surrounding primitives the GEMM program family uses, and
records them into a `Trace` when a `@T.prim_func` body is executed symbolically.
`tilelang_mesh_b200.program.recognize_gemm` then turns the trace into ONE whole-GEMM descriptor for
the B200 kernel.  Programs outside the family raise NotImplementedError — by design there is no
generic compiler and no fallback behind this facade.
"""
from __future__ import annotations

import enum
import inspect
from dataclasses import dataclass, field
from typing import Any, List, Optional, Sequence, Tuple

from .expr import BinOp, Call, Const, Expr, Var, as_affine, evaluate, is_const, wrap

__all__ = [
    "prim_func", "Tensor", "StridedTensor", "Buffer", "Kernel", "alloc_shared", "alloc_fragment", "alloc_local",
    "alloc_tmem", "alloc_barrier", "Pipelined", "Persistent", "Parallel", "serial", "grid", "unroll", "copy", "gemm", "gemm_v1",
    "gemm_v2", "clear", "fill", "use_swizzle", "annotate_layout", "mbarrier_wait_parity", "atomic_add", "ws", "barrier_wait",
    "barrier_arrive", "mbarrier_arrive", "create_list_of_mbarrier", "max", "min", "ceildiv",
    "GemmWarpPolicy", "dynamic", "symbolic", "float16", "bfloat16", "float32", "float8_e4m3", "float8_e5m2", "int32",
]

# dtype spellings accepted by the reference (tilelang/utils/tensor.py:34-51)
float16, bfloat16, float32 = "float16", "bfloat16", "float32"
float8_e4m3, float8_e5m2, int32 = "float8_e4m3", "float8_e5m2", "int32"
_DTYPE_ALIASES = {"float": "float32", "float32": "float32", "float16": "float16", "half": "float16",
                  "bfloat16": "bfloat16", "float8_e4m3": "float8_e4m3", "float8_e4m3fn": "float8_e4m3",
                  "e4m3_float8": "float8_e4m3", "float8_e5m2": "float8_e5m2", "e5m2_float8": "float8_e5m2",
                  "int32": "int32", "int8": "int8", "uint64": "uint64"}


def canon_dtype(dt) -> str:
    s = str(dt).replace("torch.", "")
    if s not in _DTYPE_ALIASES:
        raise NotImplementedError(f"dtype {dt!r} is outside the GEMM program family of this backend")
    return _DTYPE_ALIASES[s]


class GemmWarpPolicy(enum.IntEnum):
    """tilelang/primitives/gemm/base.py:13-15. Accepted and ignored: tcgen05.mma is issued by one
    thread, there is no warp partition to choose."""
    Square = 0
    FullRow = 1
    FullCol = 2

    @classmethod
    def from_warp_partition(cls, m_warp, n_warp):
        return cls.Square


# --------------------------------------------------------------------------- buffers and regions
class TensorAnnot:
    """`T.Tensor(shape, dtype)` used as a parameter annotation of a prim_func."""

    def __init__(self, shape, dtype, strides=None):
        self.shape = tuple(shape) if isinstance(shape, (tuple, list)) else (shape,)
        self.dtype = canon_dtype(dtype)
        self.strides = tuple(strides) if strides is not None else None


class _TensorFactory:
    """`T.Tensor((M, K), dtype)` and the subscript spelling `T.Tensor[(M, K), dtype]` some examples use
    (examples/warp_specialize/example_warp_specialize_gemm_softpipe_stage2.py:15-17)."""

    def __call__(self, shape, dtype="float32", *args, **kwargs):
        return TensorAnnot(shape, dtype)

    def __getitem__(self, item):
        if not isinstance(item, tuple) or len(item) < 1:
            raise TypeError("T.Tensor[shape, dtype]")
        return TensorAnnot(item[0], item[1] if len(item) > 1 else "float32")


Tensor = _TensorFactory()


def StridedTensor(shape, strides, dtype="float32", *args, **kwargs):  # noqa: N802
    return TensorAnnot(shape, dtype, strides)


Buffer = Tensor


class BufferRegion(Expr):
    """A window of a buffer: per-dimension (start, extent | None).  `extent None` = inferred later
    from the other side of a T.copy, like the reference's BufferLoad operands (copy.py:50-81).
    It is also an expression: `scales_b[bx * block_N // 128, k]` read as a scalar (a BufferLoad in the
    reference) can be multiplied and stored (example_deepgemm_fp8_2xAcc.py:70-73)."""

    def __init__(self, buffer: "TileBuffer", starts: Sequence, extents: Sequence):
        self.buffer, self.starts, self.extents = buffer, tuple(starts), tuple(extents)

    def __repr__(self):
        return f"{self.buffer.name}[{', '.join(map(repr, self.starts))}]"


class TileBuffer:
    """A global parameter tensor or an on-chip tile buffer."""

    def __init__(self, name: str, shape, dtype: str, scope: str, strides=None, param_index: Optional[int] = None):
        self.name, self.shape, self.dtype, self.scope = name, tuple(shape), dtype, scope
        self.strides, self.param_index = strides, param_index

    @property
    def is_global(self):
        return self.scope == "global"

    def _region(self, idx) -> BufferRegion:
        if not isinstance(idx, tuple):
            idx = (idx,)
        starts, extents = [], []
        for i in idx:
            if isinstance(i, slice):
                start = 0 if i.start is None else i.start
                stop = i.stop
                starts.append(start)
                if stop is None:
                    extents.append(None)
                else:
                    diff = as_affine(wrap(stop) - wrap(start))
                    if diff is None or diff[0] is not None and diff[1] != 0:
                        raise NotImplementedError("slice with a non-constant extent")
                    extents.append(diff[2])
            else:
                starts.append(i)
                extents.append(None)
        while len(starts) < len(self.shape):
            starts.append(0)
            extents.append(self.shape[len(starts) - 1])
        return BufferRegion(self, starts, extents)

    def __getitem__(self, idx):
        tr = _current_trace(optional=True)
        if tr is not None and tr.in_parallel and not self.is_global:
            return ElemLoad(self, idx if isinstance(idx, tuple) else (idx,))
        return self._region(idx)

    def __setitem__(self, idx, value):
        tr = _current_trace()
        tr.ops.append(ElemStore(self, idx if isinstance(idx, tuple) else (idx,), value, tuple(tr.parallel_vars),
                                tr.current_loop))

    def __repr__(self):
        return f"{self.name}<{self.scope} {self.shape} {self.dtype}>"


class ElemLoad(Expr):
    """Scalar element read inside a T.Parallel body (element-wise epilogue)."""
    __slots__ = ("buffer", "index")

    def __init__(self, buffer, index):
        self.buffer, self.index = buffer, index

    def __repr__(self):
        return f"{self.buffer.name}{list(self.index)}"


# --------------------------------------------------------------------------- recorded operations
@dataclass
class CopyOp:
    src: Any
    dst: Any
    loop: Optional[Var]
    coalesced_width: Optional[int] = None
    disable_tma: bool = False
    eviction_policy: Optional[str] = None


@dataclass
class GemmOp:
    A: Any
    B: Any
    C: Any
    transpose_A: bool
    transpose_B: bool
    policy: Any
    clear_accum: Any
    k_pack: int
    wg_wait: int
    mbar: Any
    loop: Optional[Var]
    variant: str = "gemm_py"


@dataclass
class ClearOp:
    buffer: TileBuffer
    value: float
    loop: Optional[Var]


@dataclass
class ElemStore:
    buffer: TileBuffer
    index: Tuple
    value: Any
    parallel_vars: Tuple
    loop: Optional[Var] = None


@dataclass
class AtomicAddOp:
    dst: Any          # BufferRegion of a global tensor (one element inside T.Parallel, or a tile origin)
    value: Any        # ElemLoad (element form) or a tile buffer (tile form)
    parallel_vars: Tuple
    loop: Optional[Var]


@dataclass
class LoopInfo:
    var: Var
    kind: str
    extent: Any
    num_stages: int = 0


@dataclass
class Trace:
    params: List[TileBuffer]
    ops: List[Any] = field(default_factory=list)
    loops: List[LoopInfo] = field(default_factory=list)
    grid: Tuple = ()
    block_vars: Tuple = ()
    threads: Any = None
    is_cpu: bool = False
    swizzle: Optional[dict] = None
    buffers: List[TileBuffer] = field(default_factory=list)
    loop_stack: List[LoopInfo] = field(default_factory=list)
    in_parallel: bool = False
    parallel_vars: List[Var] = field(default_factory=list)
    in_kernel: bool = False
    persistent: Optional[dict] = None

    @property
    def current_loop(self) -> Optional[Var]:
        for li in reversed(self.loop_stack):
            if li.kind in ("pipelined", "serial"):
                return li.var
        return None


_TRACE_STACK: List[Trace] = []


def _current_trace(optional=False) -> Optional[Trace]:
    if not _TRACE_STACK:
        if optional:
            return None
        raise RuntimeError("tile-language primitive used outside of a @T.prim_func body")
    return _TRACE_STACK[-1]


# --------------------------------------------------------------------------- prim_func
class PrimFunc:
    """What `@T.prim_func` returns: the untraced program plus its parameter signature."""

    def __init__(self, fn):
        self.orig_func = fn
        self.__name__ = getattr(fn, "__name__", "main")
        sig = inspect.signature(fn)
        self.param_names, self.annots = [], []
        for name, p in sig.parameters.items():
            ann = p.annotation
            if isinstance(ann, str):
                raise NotImplementedError(
                    "string annotations (from __future__ import annotations) hide the T.Tensor objects; "
                    "declare prim_func parameters with live T.Tensor(...) annotations")
            if not isinstance(ann, TensorAnnot):
                raise NotImplementedError(f"prim_func parameter {name!r} must be annotated with T.Tensor(shape, dtype)")
            self.param_names.append(name)
            self.annots.append(ann)
        self._trace: Optional[Trace] = None

    @property
    def params(self):
        return list(zip(self.param_names, self.annots))

    def trace(self) -> Trace:
        if self._trace is None:
            bufs = [TileBuffer(n, a.shape, a.dtype, "global", a.strides, i)
                    for i, (n, a) in enumerate(zip(self.param_names, self.annots))]
            tr = Trace(params=bufs)
            _TRACE_STACK.append(tr)
            try:
                self.orig_func(*bufs)
            finally:
                _TRACE_STACK.pop()
            self._trace = tr
        return self._trace

    def __repr__(self):
        ps = ", ".join(f"{n}: Tensor({a.shape}, {a.dtype})" for n, a in self.params)
        return f"PrimFunc {self.__name__}({ps})"


def prim_func(fn=None, **_ignored):
    if fn is None:
        return lambda f: PrimFunc(f)
    return PrimFunc(fn)


def dynamic(name: str, dtype: str = "int32"):
    """tilelang/language/__init__.py `T.dynamic` (and `tvm.te.var`, examples/dynamic_shape/example_dynamic.py:19-21):
    a shape symbol.  The B200 backend resolves a plan per concrete (M, N, K): the symbol is bound from the tensors
    of each call (jit/__init__.py) and the recognised program is cached per shape."""
    return Var(str(name), "dyn", None)


symbolic = dynamic


# --------------------------------------------------------------------------- kernel frame / loops
class Kernel:
    """`with T.Kernel(grid_x, grid_y, threads=...) as (bx, by)` — tilelang/language/kernel.py:228-232."""

    def __init__(self, *blocks, threads=None, is_cpu: bool = False, prelude=None):
        self.blocks, self.threads, self.is_cpu = blocks, threads, is_cpu

    def __enter__(self):
        tr = _current_trace()
        if tr.in_kernel:
            raise NotImplementedError("nested / multiple T.Kernel frames")
        if self.is_cpu:
            raise NotImplementedError("T.Kernel(is_cpu=True): this backend is sm_100a only and has no CPU path "
                                      "(the reference has no CPU lowering of T.gemm either: src/tl_templates/cpp/gemm.h)")
        names = ["bx", "by", "bz"]
        vars_ = tuple(Var(names[i] if i < 3 else f"b{i}", "block", b) for i, b in enumerate(self.blocks))
        tr.grid, tr.block_vars, tr.threads, tr.in_kernel = tuple(self.blocks), vars_, self.threads, True
        return vars_[0] if len(vars_) == 1 else vars_

    def __exit__(self, *exc):
        return False


def _loop(kind, extent, num_stages=0):
    tr = _current_trace()
    v = Var(f"{kind[0]}{len(tr.loops)}", "loop", extent)
    li = LoopInfo(v, kind, extent, num_stages)
    tr.loops.append(li)
    tr.loop_stack.append(li)
    try:
        yield v
    finally:
        tr.loop_stack.pop()


def Pipelined(start, stop=None, num_stages: int = 0, order=None, stage=None, sync=None, group=None):  # noqa: N802
    """tilelang/language/loop.py:57-96.  The body is traced once with a symbolic loop variable."""
    if stop is not None:
        if not is_const(start, 0):
            raise NotImplementedError("T.Pipelined with a non-zero start")
        start = stop
    return _loop("pipelined", start, int(num_stages or 0))


def serial(start, stop=None, step=None, **_):
    if stop is not None:
        if not is_const(start, 0):
            raise NotImplementedError("T.serial with a non-zero start")
        start = stop
    return _loop("serial", start)


unroll = serial


def Persistent(domain, wave_size, index, group_size=8):  # noqa: N802
    """tilelang/language/loop.py:35-54: `for bx, by in T.Persistent([tiles_m, tiles_n], sm_num, block_id)` walks the
    tile domain with `wave_size` resident blocks (examples/gemm/example_gemm_persistent.py:100-101).  The B200 kernel
    is persistent by construction, so the loop variables simply become the tile indices of the program."""
    tr = _current_trace()
    if tr.persistent is not None:
        raise NotImplementedError("more than one T.Persistent loop")
    if not any(index is bv for bv in tr.block_vars):
        raise NotImplementedError("T.Persistent: `index` must be the T.Kernel block index")
    names = ["bx", "by", "bz"]
    vs = tuple(Var("p" + names[i] if i < 3 else f"pb{i}", "block", e) for i, e in enumerate(domain))
    tr.persistent = dict(domain=tuple(domain), wave_size=wave_size, group_size=group_size)
    tr.block_vars = tuple(tr.block_vars) + vs
    yield vs[0] if len(vs) == 1 else vs


def Parallel(*extents, coalesced_width=None):  # noqa: N802
    """tilelang/language/loop.py:12.  Recognised as an element-wise epilogue over the accumulator tile, or as an
    element-wise global->shared staging loop (examples/gemm/example_gemm_schedule.py:21-22)."""
    tr = _current_trace()
    vs = tuple(Var(f"p{i}", "parallel", e) for i, e in enumerate(extents))
    tr.in_parallel, tr.parallel_vars = True, list(vs)
    try:
        yield vs[0] if len(vs) == 1 else vs
    finally:
        tr.in_parallel, tr.parallel_vars = False, []


def grid(*extents):
    raise NotImplementedError("T.grid scalar loops (the reference's CPU-target matmul form) are not a T.gemm program; "
                              "this backend accelerates the T.gemm / T.copy path only")


# --------------------------------------------------------------------------- allocation
def _alloc(shape, dtype, scope, prefix):
    tr = _current_trace()
    shape = tuple(shape) if isinstance(shape, (tuple, list)) else (shape,)
    b = TileBuffer(f"{prefix}{len(tr.buffers)}", shape, canon_dtype(dtype), scope)
    tr.buffers.append(b)
    return b


def alloc_shared(shape, dtype, scope="shared.dyn"):
    return _alloc(shape, dtype, scope, "shared")


def alloc_fragment(shape, dtype, scope="local.fragment"):
    return _alloc(shape, dtype, scope, "frag")


def alloc_local(shape, dtype, scope="local"):
    return _alloc(shape, dtype, scope, "local")


def alloc_tmem(shape, dtype):
    """tilelang/language/allocate.py:172-198: 2-D accumulator in tensor memory."""
    return _alloc(shape, dtype, "shared.tmem", "tmem")


def alloc_barrier(arrive_count):
    """tilelang/language/allocate.py:160: user-managed mbarrier.  The kernel owns its own barriers, so
    the handle is only carried through to T.gemm(mbar=...) / T.mbarrier_wait_parity."""
    return _alloc((1,), "uint64", "shared.barrier", "mbar")


# --------------------------------------------------------------------------- the two tile operators
def copy(src, dst, coalesced_width: Optional[int] = None, disable_tma: bool = False,
         eviction_policy: Optional[str] = None):
    """T.copy — signature of tilelang/language/copy.py:13-17.

    `coalesced_width`, `disable_tma` and `eviction_policy` are accepted for compatibility; staging
    is always TMA bulk-tensor loads into 128B-swizzled shared memory on this backend."""
    tr = _current_trace()
    if eviction_policy not in (None, "evict_normal", "evict_first", "evict_last"):
        raise ValueError(f"T.copy: unknown eviction_policy {eviction_policy!r}")
    if isinstance(src, TileBuffer) and isinstance(dst, TileBuffer) and src.shape != dst.shape:
        raise ValueError(f"T.copy: shape mismatch {src.shape} vs {dst.shape}")  # copy.py:46-47
    tr.ops.append(CopyOp(src, dst, tr.current_loop, coalesced_width, bool(disable_tma), eviction_policy))


def _gemm_impl(variant, A, B, C, transpose_A, transpose_B, policy, clear_accum, k_pack, wg_wait, mbar):
    tr = _current_trace()
    for name, x in (("A", A), ("B", B), ("C", C)):
        if not isinstance(x, (TileBuffer, BufferRegion)):
            raise TypeError(f"T.gemm: operand {name} must be a buffer or buffer region")
    a_shape = (A.buffer.shape if isinstance(A, BufferRegion) else A.shape)[-2:]
    b_shape = (B.buffer.shape if isinstance(B, BufferRegion) else B.shape)[-2:]
    K_A = a_shape[0] if transpose_A else a_shape[1]
    K_B = b_shape[1] if transpose_B else b_shape[0]
    if K_A != K_B:  # tilelang/language/gemm.py:78-81
        raise AssertionError(f"T.gemm K shape check failed: K_A = {K_A}, K_B = {K_B}")
    tr.ops.append(GemmOp(A, B, C, bool(transpose_A), bool(transpose_B), policy, clear_accum, int(k_pack), int(wg_wait),
                         mbar, tr.current_loop, variant))


def gemm_v1(A, B, C, transpose_A: bool = False, transpose_B: bool = False, policy=GemmWarpPolicy.Square,
            clear_accum=False, k_pack: int = 1, wg_wait: int = 0, mbar=None):
    """tilelang/language/gemm.py:105-132 (op tl.tileop.gemm)."""
    return _gemm_impl("gemm", A, B, C, transpose_A, transpose_B, policy, clear_accum, k_pack, wg_wait, mbar)


def gemm_v2(A, B, C, transpose_A: bool = False, transpose_B: bool = False, policy=GemmWarpPolicy.Square,
            clear_accum=False, k_pack: int = 1, wg_wait: int = 0, mbar=None):
    """tilelang/language/gemm.py:136-160 (op tl.tileop.gemm_py) — the default `T.gemm` (:163)."""
    return _gemm_impl("gemm_py", A, B, C, transpose_A, transpose_B, policy, clear_accum, k_pack, wg_wait, mbar)


gemm = gemm_v2


# --------------------------------------------------------------------------- small primitives
def clear(buffer):
    """tilelang/language/fill.py:39."""
    tr = _current_trace()
    tr.ops.append(ClearOp(buffer, 0.0, tr.current_loop))


def fill(buffer, value):
    tr = _current_trace()
    tr.ops.append(ClearOp(buffer, float(value), tr.current_loop))


def atomic_add(dst, value, memory_order=None, return_prev: bool = False, use_tma: bool = False):
    """tilelang/language/atomic.py:115-119.  Recognised only as the output step of a split-K GEMM program
    (examples/gemm_splitk/example_tilelang_gemm_splitk.py:39-40): partial tiles added into a global C."""
    if return_prev:
        raise NotImplementedError("T.atomic_add(return_prev=True) is outside the GEMM program family")
    tr = _current_trace()
    tr.ops.append(AtomicAddOp(dst, value, tuple(tr.parallel_vars), tr.current_loop))


def use_swizzle(panel_size: int = 10, order: str = "row", enable: bool = True):
    """tilelang/language/annotations.py:16-21: L2 rasterisation hint -> persistent scheduler's raster group."""
    _current_trace().swizzle = dict(panel_size=panel_size, order=order, enable=bool(enable))


def annotate_layout(layout_map):
    """tilelang/language/annotations.py:24-35.  Shared-memory layouts are fixed (128B swizzle) here."""
    return None


def mbarrier_wait_parity(mbarrier, parity):
    """tilelang/language/builtin.py:188.  The kernel's own pipeline owns MMA completion."""
    return None


class ws:  # noqa: N801 - reference spelling `with T.ws(0):`
    """tilelang/language/warpgroup.py: statements for one warp group of a hand-specialised kernel
    (examples/warp_specialize/example_warp_specialize_gemm_copy_0_gemm_1.py:32-48).  The B200 kernel has its own
    fixed producer / MMA / epilogue roles, so the annotation only groups statements here."""

    def __init__(self, *warp_group_ids):
        self.ids = warp_group_ids

    def __enter__(self):
        _current_trace()
        return self

    def __exit__(self, *exc):
        return False


def barrier_wait(barrier_id, parity=None):
    """tilelang/language/builtin.py (alias of mbarrier_wait_parity): the kernel owns its pipeline barriers."""
    return None


def barrier_arrive(barrier_id):
    return None


def mbarrier_arrive(mbarrier):
    return None


def create_list_of_mbarrier(*args):
    return None


def max(a, b):  # noqa: A001 - reference spelling T.max
    return Call("max", (a, b))


def min(a, b):  # noqa: A001
    return Call("min", (a, b))


def ceildiv(a, b):
    if isinstance(a, int) and isinstance(b, int):
        return -(-a // b)
    return (wrap(a) + (wrap(b) - 1)) // wrap(b)
