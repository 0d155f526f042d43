"""Binding into a genuine TileLang: from a real ``tvm.tir.PrimFunc`` to a running B200 kernel.

Everything else in this package recognises programs traced by its own facade (``language/``).  This module
works on what the reference itself builds: the TIR of a ``@T.prim_func`` whose body holds the tile-operator calls
``tl.tileop.copy`` / ``tl.tileop.gemm(_py)`` / ``tl.tileop.fill`` with ``tl.tileop.region`` operands
(tilelang/language/gemm.py:99-101, copy.py, fill.py; decoded by the C++ operators in src/op/gemm_py.cc:53-93,
src/op/copy.cc).  It needs an importable ``tilelang`` that is the real compiler (the wheel of the image, or the
reference once built) - never this repo's ``compat/tilelang`` facade.

    decode_gemm_call(call)        the 19 positional arguments of the gemm tile operator, as GemmPy::GemmPy reads them
    trace_from_primfunc(func)     the TIR body re-expressed as this backend's trace (same records the facade makes)
    recognize_primfunc(func)      -> program.GemmProgram, by the same recogniser the facade uses
    wrapper_source(prog)          C source of a per-program host library exporting the reference's generated trio
                                  `init` / `call(A, B, C, stream)` / `get_last_error` (jit/adapter/wrapper.py:31-51)
    build_wrapper_library(...)    compiles it (gcc, links libtlb200.so)
    compile(func, out_idx)        -> tilelang.jit.kernel.JITKernel.from_database(kernel_lib_path=...)
                                  (tilelang/jit/kernel.py:143-183): the reference's own adapter then drives the kernel

`call` has no workspace argument and the framework above it owns device memory, so it forwards to
`tlb200_gemm_static_scratch` (include/tlb200.h).
"""
from __future__ import annotations

import hashlib
import os
import subprocess
import tempfile
from typing import Any, Dict, List, Optional

from . import _abi
from . import language as L
from .language.expr import BinOp, Call, Const, Var
from .program import GemmProgram, NotAGemmProgram, recognize_gemm

GEMM_OPS = ("tl.tileop.gemm", "tl.tileop.gemm_py", "tl.gemm", "tl.gemm_py")
COPY_OPS = ("tl.tileop.copy", "tl.copy")
FILL_OPS = ("tl.tileop.fill", "tl.fill")
REGION_OPS = ("tl.tileop.region", "tl.region")
# calls that order or annotate but carry no data (the kernel has its own barriers and layouts)
IGNORED_OPS = ("tl.mbarrier_wait_parity", "tl.mbarrier_arrive", "tl.mbarrier_expect_tx", "tl.ptx_arrive_barrier",
               "tl.sync_grid", "tl.no_set_max_nreg", "tl.set_max_nreg", "tl.annotate_layout", "tl.sync_threads",
               "tir.tvm_storage_sync")


def _tvm():
    try:
        from tilelang import tvm  # the compiler's own TVM
    except Exception as e:  # noqa: BLE001
        raise RuntimeError("tir_frontend needs a real TileLang (`from tilelang import tvm`): " + repr(e)) from None
    import tilelang
    if os.path.basename(os.path.dirname(os.path.dirname(os.path.abspath(tilelang.__file__)))) == "compat":
        raise RuntimeError("`tilelang` resolves to this repo's compat/ facade; tir_frontend binds into the real compiler")
    return tvm


class UnknownOp:
    """Anything in the body this front end does not understand: the recogniser's sweep rejects the program."""

    def __init__(self, what, loop):
        self.what, self.loop = what, loop

    def __repr__(self):
        return f"UnknownOp({self.what})"


def decode_gemm_call(call) -> Dict[str, Any]:
    """The 19 positional arguments of `tl.tileop.gemm` / `gemm_py` (tilelang/language/gemm.py:99-101), named as
    GemmPy::GemmPy reads them (src/op/gemm_py.cc:53-93)."""
    a = list(call.args)
    if len(a) < 14:
        raise NotAGemmProgram(f"gemm tile operator with {len(a)} arguments (expected 14-19)")
    names = ["A", "B", "C", "trans_A", "trans_B", "M", "N", "K", "policy", "clear_accum", "stride_A", "stride_B",
             "offset_A", "offset_B", "k_pack", "wg_wait", "mbar", "c_coord0", "c_coord1"]
    out = dict(zip(names, a))
    out.setdefault("k_pack", 1)
    out.setdefault("wg_wait", 0)
    out.setdefault("mbar", None)
    return out


class _Converter:
    def __init__(self, func):
        self.tvm = _tvm()
        self.tir = self.tvm.tir
        self.func = func
        self.vars: Dict[Any, Var] = {}
        self.bufs: Dict[Any, L.TileBuffer] = {}
        self.trace = L.Trace(params=[])
        self.loop_stack: List[L.LoopInfo] = []
        self.parallel: List[Var] = []

    # ---- expressions
    def _int(self, e):
        tir = self.tir
        if isinstance(e, (int, bool)):
            return int(e)
        if isinstance(e, tir.IntImm):
            return int(e.value)
        return None

    def expr(self, e):
        tir = self.tir
        if isinstance(e, (int, float, bool)):
            return e
        if isinstance(e, tir.IntImm):
            return bool(e.value) if str(e.dtype) == "bool" else int(e.value)
        if isinstance(e, tir.FloatImm):
            return float(e.value)
        if isinstance(e, tir.Var):
            v = self.vars.get(e)
            if v is None:  # a dynamic shape symbol
                v = self.vars[e] = Var(str(e.name), "dyn", None)
            return v
        if isinstance(e, tir.Cast):
            return self.expr(e.value)
        binops = {tir.Add: "+", tir.Sub: "-", tir.Mul: "*", tir.FloorDiv: "//", tir.FloorMod: "%", tir.EQ: "==",
                  tir.NE: "!=", tir.LT: "<", tir.LE: "<=", tir.GT: ">", tir.GE: ">=", tir.Div: "//", tir.Mod: "%"}
        for cls, op in binops.items():
            if isinstance(e, cls):
                a, b = self.expr(e.a), self.expr(e.b)
                if not isinstance(a, L.Expr) and not isinstance(b, L.Expr):
                    return _fold(op, a, b)
                return BinOp(op, _wrap(a), _wrap(b))
        if isinstance(e, tir.Max):
            return Call("max", [self.expr(e.a), self.expr(e.b)])
        if isinstance(e, tir.Min):
            return Call("min", [self.expr(e.a), self.expr(e.b)])
        if isinstance(e, tir.BufferLoad):
            buf = self.buffer(e.buffer)
            idx = tuple(self.expr(i) for i in e.indices)
            if buf.is_global:
                return L.BufferRegion(buf, idx, [None] * len(idx))
            return L.ElemLoad(buf, idx)
        raise NotAGemmProgram(f"expression {e} ({type(e).__name__}) is outside what the GEMM family uses")

    # ---- buffers and regions
    def buffer(self, b) -> L.TileBuffer:
        tb = self.bufs.get(b)
        if tb is None:
            shape = tuple(self._int(s) if self._int(s) is not None else self.expr(s) for s in b.shape)
            tb = self.bufs[b] = L.TileBuffer(str(b.name), shape, L.canon_dtype(str(b.dtype)), str(b.scope()) or "global")
            self.trace.buffers.append(tb)
        return tb

    def region(self, arg):
        """`tl.tileop.region(BufferLoad, access_mask, extent...)` -> a window of the buffer; a whole on-chip tile is
        passed as the tile buffer itself (what the facade records for `T.copy(A[...], A_shared)`)."""
        tir = self.tir
        if isinstance(arg, tir.Call) and str(arg.op.name) in REGION_OPS:
            load, extents = arg.args[0], [self._int(x) for x in arg.args[2:]]
            buf = self.buffer(load.buffer)
            starts = [self.expr(i) for i in load.indices]
            ext = list(extents[-len(starts):]) if len(extents) >= len(starts) else [None] * len(starts)
            whole = (not buf.is_global and all(self._int_or(s) == 0 for s in starts) and tuple(ext) == tuple(buf.shape))
            return buf if whole else L.BufferRegion(buf, starts, ext)
        if isinstance(arg, tir.BufferLoad):
            buf = self.buffer(arg.buffer)
            return L.BufferRegion(buf, [self.expr(i) for i in arg.indices], [None] * len(arg.indices))
        raise NotAGemmProgram(f"tile-operator operand {arg} is not a tl.region / BufferLoad")

    def _int_or(self, x):
        return x if isinstance(x, int) else (x.value if isinstance(x, Const) else None)

    # ---- statements
    @property
    def k_loop(self) -> Optional[Var]:
        for li in reversed(self.loop_stack):
            if li.kind in ("pipelined", "serial"):
                return li.var
        return None

    def stmt(self, s):
        tir, tr = self.tir, self.trace
        if isinstance(s, tir.SeqStmt):
            for x in s.seq:
                self.stmt(x)
        elif isinstance(s, tir.BlockRealize):
            self.stmt(s.block)
        elif isinstance(s, tir.Block):
            for b in s.alloc_buffers:
                self.buffer(b)
            if s.init is not None:
                tr.ops.append(UnknownOp("block init", self.k_loop))
            self.stmt(s.body)
        elif isinstance(s, tir.AttrStmt):
            key = str(s.attr_key)
            if key == "thread_extent":
                tag, ext = str(s.node.thread_tag), self._int(s.value)
                if tag.startswith("blockIdx"):
                    v = Var({"blockIdx.x": "bx", "blockIdx.y": "by", "blockIdx.z": "bz"}[tag], "block", ext)
                    self.vars[s.node.var] = v
                    tr.block_vars = tuple(tr.block_vars) + (v,)
                    tr.grid = tuple(tr.grid) + (ext,)
                    tr.in_kernel = True
                elif tag == "threadIdx.x":
                    tr.threads = ext
                    self.vars[s.node.var] = Var("tx", "thread", ext)
                else:
                    self.vars[s.node.var] = Var(tag, "thread", ext)
            elif key == "threadblock_swizzle_pattern":
                tr.swizzle = dict(pattern=str(s.value))
            self.stmt(s.body)
        elif isinstance(s, tir.For):
            if self._int(s.min) != 0:
                raise NotAGemmProgram(f"loop {s.loop_var} does not start at 0")
            ext = self._int(s.extent) if self._int(s.extent) is not None else self.expr(s.extent)
            if int(s.kind) == int(tir.ForKind.PARALLEL):
                v = Var(str(s.loop_var.name), "parallel", ext)
                self.vars[s.loop_var] = v
                self.parallel.append(v)
                try:
                    self.stmt(s.body)
                finally:
                    self.parallel.pop()
                return
            ann = {str(k): v for k, v in dict(s.annotations).items()}
            kind = "pipelined" if "num_stages" in ann else "serial"
            v = Var(str(s.loop_var.name), "loop", ext)
            self.vars[s.loop_var] = v
            li = L.LoopInfo(v, kind, ext, self._int(ann.get("num_stages", 0)) or 0)
            tr.loops.append(li)
            self.loop_stack.append(li)
            try:
                self.stmt(s.body)
            finally:
                self.loop_stack.pop()
        elif isinstance(s, tir.Evaluate):
            self.call(s.value)
        elif isinstance(s, tir.BufferStore):
            if not self.parallel:
                tr.ops.append(UnknownOp(f"scalar store to {s.buffer.name}", self.k_loop))
                return
            buf = self.buffer(s.buffer)
            idx = tuple(self.expr(i) for i in s.indices)
            tr.ops.append(L.ElemStore(buf, idx, _wrap(self.expr(s.value)), tuple(self.parallel), self.k_loop))
        elif isinstance(s, tir.LetStmt):
            self.vars[s.var] = _wrap(self.expr(s.value))
            self.stmt(s.body)
        else:
            tr.ops.append(UnknownOp(type(s).__name__, self.k_loop))

    def call(self, c):
        tir, tr = self.tir, self.trace
        if not isinstance(c, tir.Call):
            return
        name = str(getattr(c.op, "name", c.op))
        if name in COPY_OPS:
            tr.ops.append(L.CopyOp(self.region(c.args[0]), self.region(c.args[1]), self.k_loop))
        elif name in GEMM_OPS:
            g = decode_gemm_call(c)
            ca = self.expr(g["clear_accum"])
            mbar = g["mbar"]
            tr.ops.append(L.GemmOp(self.region(g["A"]), self.region(g["B"]), self.region(g["C"]),
                                   bool(self._int(g["trans_A"])), bool(self._int(g["trans_B"])), self._int(g["policy"]),
                                   ca, self._int(g["k_pack"]) or 1, self._int(g["wg_wait"]) or 0,
                                   None if self._int(mbar) == 0 else mbar, self.k_loop,
                                   variant="gemm_py" if name.endswith("gemm_py") else "gemm"))
        elif name in FILL_OPS:
            dst = self.region(c.args[0])
            val = self.expr(c.args[1])
            if isinstance(dst, L.BufferRegion) or isinstance(val, L.Expr):
                tr.ops.append(UnknownOp("partial or symbolic T.fill", self.k_loop))
            else:
                tr.ops.append(L.ClearOp(dst, float(val), self.k_loop))
        elif name in IGNORED_OPS:
            pass
        else:
            tr.ops.append(UnknownOp(name, self.k_loop))

    def run(self) -> L.Trace:
        func = self.func
        for i, p in enumerate(func.params):
            b = func.buffer_map.get(p)
            if b is None:
                raise NotAGemmProgram(f"parameter {p} is not a tensor")
            tb = self.buffer(b)
            tb.scope, tb.param_index = "global", i
            tb.strides = tuple(self._int(s) for s in b.strides) if len(b.strides) else None
            self.trace.params.append(tb)
        self.stmt(func.body)
        return self.trace


def _wrap(x):
    from .language.expr import wrap
    return wrap(x)


def _fold(op, a, b):
    return {"+": lambda: a + b, "-": lambda: a - b, "*": lambda: a * b, "//": lambda: a // b, "%": lambda: a % b,
            "==": lambda: a == b, "!=": lambda: a != b, "<": lambda: a < b, "<=": lambda: a <= b, ">": lambda: a > b,
            ">=": lambda: a >= b}[op]()


def trace_from_primfunc(func) -> "L.Trace":
    """The TIR body of a real `@T.prim_func` as this backend's trace records."""
    return _Converter(func).run()


class _TracedFunc:
    """What `program.recognize_gemm` expects of a function: `.trace()` and `.params`."""

    def __init__(self, trace):
        self._trace = trace
        self.params = [(b.name, b) for b in trace.params]

    def trace(self):
        return self._trace


def recognize_primfunc(func) -> GemmProgram:
    tr = trace_from_primfunc(func)
    for b in tr.params:
        if any(not isinstance(s, int) for s in b.shape):
            raise NotAGemmProgram("dynamic shapes are bound per call by this repo's own JIT; the spliced wrapper library is "
                                  "generated for static shapes")
        if b.strides is not None and len(b.shape) == 2 and (b.strides[1] != 1 or b.strides[0] != b.shape[1]):
            raise NotAGemmProgram(f"parameter {b.name} is not contiguous (strides {b.strides})")
    return recognize_gemm(_TracedFunc(tr))


# ------------------------------------------------------------------------------------------ the wrapper library
_DT = {"float16": "TLB200_FLOAT16", "bfloat16": "TLB200_BFLOAT16", "float8_e4m3": "TLB200_FLOAT8_E4M3",
       "float8_e5m2": "TLB200_FLOAT8_E5M2", "float32": "TLB200_FLOAT32", "int8": "TLB200_INT8", "int32": "TLB200_INT32"}


def wrapper_source(p: GemmProgram) -> str:
    """Host library of ONE program with the reference's generated C ABI (wrapper.py:31-51):
    `int init()`, `int call(<tensor pointers in prim_func order>, cudaStream_t)`, `const char* get_last_error()`."""
    order = sorted([(p.a_param, "a"), (p.b_param, "b"), (p.c_param, "c")])
    args = ", ".join(f"void* arg{i}" for i, _ in order)
    pick = {nm: f"arg{i}" for i, nm in order}
    lda = p.M if p.trans_A else p.K
    ldb = p.K if p.trans_B else p.N
    epi = ("TLB200_EPI_RELU" if p.relu else "TLB200_EPI_NONE") + (" | TLB200_EPI_ACCUMULATE" if p.accumulate else "")
    return f'''// Generated by tilelang_mesh_b200.tir_frontend for: {p.summary()}
// The reference generates this trio per kernel (tilelang/jit/adapter/wrapper.py:31-51, 202-295); here `call`
// describes the whole contraction to libtlb200.so instead of launching generated device code.
#include <string.h>
#include "tlb200.h"

static tlb200_gemm_desc g_desc;
static char g_err[1024];

const char* get_last_error(void) {{ return g_err; }}

int init(void) {{
  g_err[0] = 0;
  if (tlb200_abi_version() != TLB200_ABI_VERSION) {{
    strncpy(g_err, "libtlb200.so speaks another ABI version than this wrapper was generated for", sizeof(g_err) - 1);
    return -1;
  }}
  memset(&g_desc, 0, sizeof(g_desc));
  g_desc.abi_version = TLB200_ABI_VERSION;
  g_desc.m = {p.M}; g_desc.n = {p.N}; g_desc.k = {p.K};
  g_desc.ab_dtype = {_DT[p.in_dtype]}; g_desc.c_dtype = {_DT[p.out_dtype]};
  g_desc.trans_a = {int(p.trans_A)}; g_desc.trans_b = {int(p.trans_B)};
  g_desc.lda = {lda}; g_desc.ldb = {ldb}; g_desc.ldc = {p.N};
  g_desc.alpha = {float(p.alpha)!r}f;
  g_desc.epilogue = {epi};
  return 0;
}}

int call({args}, void* stream) {{
  // no workspace argument in this signature and the caller owns device memory: split-K scratch is the library's own
  int rc = tlb200_gemm_static_scratch(&g_desc, {pick["a"]}, {pick["b"]}, {pick["c"]}, stream);
  if (rc != 0) {{
    strncpy(g_err, tlb200_last_error(), sizeof(g_err) - 1);
    g_err[sizeof(g_err) - 1] = 0;
  }}
  return rc;
}}
'''


def build_wrapper_library(p: GemmProgram, out_dir: Optional[str] = None) -> str:
    src = wrapper_source(p)
    tag = hashlib.sha256(src.encode()).hexdigest()[:16]
    out_dir = out_dir or os.path.join(tempfile.gettempdir(), "tlb200_wrappers")
    os.makedirs(out_dir, exist_ok=True)
    c_path, so_path = os.path.join(out_dir, f"gemm_{tag}.c"), os.path.join(out_dir, f"gemm_{tag}.so")
    if os.path.exists(so_path):
        return so_path
    with open(c_path, "w") as f:
        f.write(src)
    lib_dir = os.path.dirname(_abi.LIB_PATH)
    inc = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include")
    cmd = ["gcc", "-O2", "-shared", "-fPIC", "-I", inc, c_path, "-o", so_path + ".tmp", "-L", lib_dir, "-ltlb200",
           "-Wl,-rpath," + lib_dir]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("gcc failed on the wrapper library:\n" + res.stdout + res.stderr)
    os.replace(so_path + ".tmp", so_path)
    return so_path


def compile(func, out_idx=None, out_dir: Optional[str] = None, execution_backend: str = "cython"):  # noqa: A001
    """A genuine `tilelang.jit.kernel.JITKernel` whose library is this backend: the program's TIR is recognised, the
    wrapper library built, and the kernel created with `JITKernel.from_database(kernel_lib_path=...)`
    (tilelang/jit/kernel.py:143-183) so that the compiler's own adapter marshals tensors and calls `call`."""
    _tvm()
    import inspect

    from tilelang.engine.param import KernelParam
    from tilelang.jit.kernel import JITKernel
    prog = recognize_primfunc(func)
    lib_path = build_wrapper_library(prog, out_dir)
    params = [KernelParam.from_buffer(func.buffer_map[p]) for p in func.params]
    src = wrapper_source(prog)
    kw = dict(func=func, kernel_lib_path=lib_path, params=params, target="cuda", target_host="c",
              out_idx=out_idx if out_idx is not None else [prog.c_param], execution_backend=execution_backend,
              pass_configs={})
    sig = inspect.signature(JITKernel.from_database).parameters
    if "host_kernel_source" in sig:       # TileLang >= 0.1.8
        kw.update(host_kernel_source=src, device_kernel_source="// device code: libtlb200.so (gemm_tcgen05_kernel)\n")
    else:                                  # the reference snapshot (0.1.7): one source argument
        kw.update(kernel_global_source=src)
    if "execution_backend" in sig and execution_backend not in str(sig["execution_backend"].annotation):
        kw["execution_backend"] = "ctypes" if "ctypes" in str(sig["execution_backend"].annotation) else execution_backend
    kernel = JITKernel.from_database(**kw)
    kernel.tlb200_program = prog
    return kernel
