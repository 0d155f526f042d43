"""Recognise the GEMM program family in a traced tile program.

The reference lowers `T.copy` / `T.gemm` tile operators one by one (src/op/copy.cc:868-901,
src/op/gemm.cc:437-573, src/op/gemm_py.cc:245-282) and relies on ~48 generic passes to stitch
pipelining, barriers and layouts around them.  The B200 kernel embodies all of that, so the
"lowering" here is pattern recognition: the trace of

    with T.Kernel(ceildiv(N, bN), ceildiv(M, bM), threads=...) as (bx, by):
        [T.clear(C_acc)]
        for k in T.Pipelined(ceildiv(K, bK), num_stages=...):
            T.copy(A[by*bM, k*bK], A_shared); T.copy(B[...], B_shared)
            T.gemm(A_shared, B_shared, C_acc, transpose_A, transpose_B, [clear_accum=k==0], ...)
        [element-wise epilogue over C_acc]
        T.copy(C_acc, [C_local ->] [C_shared ->] C[by*bM, bx*bN])

is turned into ONE whole-problem descriptor.  Anything that does not match raises
NotImplementedError naming the first thing that fell outside the family.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Optional

from . import language as T
from .language.expr import BinOp, Call, Const, Expr, Var, as_affine, as_linear, is_const


@dataclass
class GemmProgram:
    M: int
    N: int
    K: int
    in_dtype: str
    out_dtype: str
    accum_dtype: str
    trans_A: bool
    trans_B: bool
    a_param: int  # positions of A, B, C in the prim_func signature
    b_param: int
    c_param: int
    block_M: int
    block_N: int
    block_K: int
    num_stages: int
    threads: Optional[int]
    relu: bool = False
    alpha: float = 1.0
    accumulate: bool = False       # C += A B (split-K program that adds its partial tiles into C with T.atomic_add)
    split_k: int = 1               # K slices of the program's grid (a hint: the backend picks its own split)
    accumulator: str = "fragment"  # "fragment" (register accumulators in the reference) or "tmem"
    rasterization: Optional[dict] = None
    notes: List[str] = field(default_factory=list)

    def summary(self) -> str:
        lay = ("T" if self.trans_A else "N") + ("T" if self.trans_B else "N")
        return (f"gemm {self.M}x{self.N}x{self.K} {self.in_dtype}->{self.out_dtype} (acc {self.accum_dtype}) {lay} "
                f"tile-hint {self.block_M}x{self.block_N}x{self.block_K} stages {self.num_stages}"
                + (" +relu" if self.relu else "") + (f" alpha={self.alpha}" if self.alpha != 1.0 else ""))


class NotAGemmProgram(NotImplementedError):
    pass


def _fail(msg):
    raise NotAGemmProgram("this tile program is outside the GEMM family the B200 backend accelerates: " + msg +
                          " (there is no generic compiler or fallback behind this facade)")


def _buf(x):
    return x.buffer if isinstance(x, T.BufferRegion) else x


def _tile_origin(region, want_vars: Dict[str, Var], what: str):
    """For a global-tensor window return [(var, scale, offset)] per dimension."""
    if isinstance(region, T.TileBuffer):
        return [(None, 0, 0)] * len(region.shape)
    out = []
    for s in region.starts:
        aff = as_affine(s)
        if aff is None:
            lin = as_linear(s)  # split-K programs index K with two variables: bz*(K/split) + ko*block_K
            if lin is None or lin[1] != 0 or len(lin[0]) != 2:
                _fail(f"{what}: index {s!r} is not of the form var*block + const")
            aff = (lin[0], None, 0)
        out.append(aff)
    return out


def _elementwise_staging(o, nm):
    """`for i, j in T.Parallel(bR, bC): X_shared[i, j] = X[r0 + i, c0 + j]` is a staging copy written element by
    element (examples/gemm/example_gemm_schedule.py:21-22); rewrite it as the T.copy it means."""
    tile, src = o.buffer, o.value
    pv = tuple(o.parallel_vars)
    if len(o.index) != len(pv) or any(i is not v for i, v in zip(o.index, pv)):
        _fail(f"element-wise fill of {nm}_shared must index the tile with the T.Parallel variables in order")
    if tuple(v.extent for v in pv) != tuple(tile.shape):
        _fail(f"element-wise fill of {nm}_shared does not cover the tile {tile.shape}")
    if len(src.starts) != len(pv):
        _fail(f"element-wise fill of {nm}_shared reads a {len(src.starts)}-D element of {src.buffer.name}")
    starts = []
    for d, s in enumerate(src.starts):
        lin = as_linear(s)
        if lin is None:
            _fail(f"element-wise fill of {nm}_shared: index {s!r} is not linear")
        terms, off = lin
        if terms.pop(pv[d], 0) != 1 or any(v in terms for v in pv):
            _fail(f"element-wise fill of {nm}_shared: dimension {d} of {src.buffer.name} must be origin + "
                  f"{pv[d]!r} (got {s!r})")
        if len(terms) > 1:
            _fail(f"element-wise fill of {nm}_shared: origin {s!r} mixes several indices")
        base = off
        for v, c in terms.items():
            base = v * c + off
        starts.append(base)
    return T.CopyOp(T.BufferRegion(src.buffer, starts, tile.shape), tile, o.loop)


def recognize_gemm(func: "T.PrimFunc", env: Optional[dict] = None) -> GemmProgram:
    """`env` binds dynamic shape symbols (T.dynamic / tvm.te.var) to the sizes of the tensors of one call."""
    from .language.expr import evaluate
    env = env or {}

    def cv(x):  # concrete value of a shape entry / extent under `env`
        if isinstance(x, Expr):
            try:
                return int(evaluate(x, env))
            except KeyError:
                _fail(f"shape symbol in {x!r} is not bound (dynamic shapes are resolved per call from the tensors)")
        return x

    tr = func.trace()
    if not tr.in_kernel:
        _fail("no T.Kernel frame")
    gemms = [o for o in tr.ops if isinstance(o, T.GemmOp)]
    if len(gemms) != 1:
        _fail(f"expected exactly one T.gemm, found {len(gemms)}")
    g = gemms[0]
    if g.loop is None:
        _fail("T.gemm is not inside a T.Pipelined / T.serial K loop")
    kloop = next(li for li in tr.loops if li.var is g.loop)
    A_s, B_s, C_acc = _buf(g.A), _buf(g.B), _buf(g.C)
    notes = []
    used = [g]  # ops the pattern accounts for; every other recorded op makes the program something else (swept at the end)

    def through_fragment(b, nm):
        # gemm_rs / gemm_sr / gemm_rr forms of the reference's tile-library tests (matmul_rs: T.copy(A_shared, A_frag);
        # T.gemm(A_frag, ...), testing/python/tilelibrary/test_tilelang_tilelibrary_gemm.py:137-186): the register copy
        # of an operand tile is the same tile, so follow it back to the shared buffer that was filled from global.
        if b.scope != "local.fragment":
            return b
        cs = [o for o in tr.ops if isinstance(o, T.CopyOp) and _buf(o.dst) is b]
        if len(cs) != 1 or cs[0].loop is not g.loop or isinstance(cs[0].src, T.BufferRegion) or \
                _buf(cs[0].src).is_global or tuple(_buf(cs[0].src).shape) != tuple(b.shape) or _buf(cs[0].src).dtype != b.dtype:
            _fail(f"T.gemm operand {nm} is a fragment that is not a plain copy of a shared tile filled in the same K loop")
        notes.append(f"operand {nm} read from a fragment copy of its shared tile (gemm_{'r' if nm == 'A' else 's'}* form): "
                     "tcgen05.mma reads the shared tile directly")
        used.append(cs[0])
        return _buf(cs[0].src)

    A_s, B_s = through_fragment(A_s, "A"), through_fragment(B_s, "B")
    for nm, b in (("A", A_s), ("B", B_s)):
        if b.is_global or not b.scope.startswith("shared") or b.scope in ("shared.tmem", "shared.barrier"):
            _fail(f"T.gemm operand {nm} must be a shared-memory tile (gemm_ss); got scope {b.scope!r} "
                  "(reference tcgen05 path supports gemm_ss only: tileop/gemm/gemm_tcgen05.py:84-86)")
    if C_acc.scope not in ("local.fragment", "shared.tmem"):
        _fail(f"accumulator scope {C_acc.scope!r}")
    if g.k_pack not in (1, 2):
        raise ValueError("T.gemm: k_pack must be 1 or 2")  # src/op/gemm.cc:84-86
    if len(A_s.shape) != 2 or len(B_s.shape) != 2 or len(C_acc.shape) != 2:
        _fail("tile buffers must be 2-D")

    bM, bN = C_acc.shape
    bK = A_s.shape[0] if g.transpose_A else A_s.shape[1]
    exp_a = (bK, bM) if g.transpose_A else (bM, bK)
    exp_b = (bN, bK) if g.transpose_B else (bK, bN)
    if tuple(A_s.shape) != exp_a or tuple(B_s.shape) != exp_b:
        _fail(f"tile shapes A{A_s.shape} B{B_s.shape} C{C_acc.shape} are inconsistent with transpose flags")
    if A_s.dtype != B_s.dtype:
        _fail(f"A is {A_s.dtype} but B is {B_s.dtype}")

    # ---- staging copies: global -> A_shared / B_shared inside the K loop
    def staging(tile, nm):
        cs = [o for o in tr.ops if isinstance(o, T.CopyOp) and _buf(o.dst) is tile]
        raw = list(cs)
        for o in tr.ops:
            if isinstance(o, T.ElemStore) and o.buffer is tile:
                # any element-wise write to an operand tile other than the staging fill itself (e.g. scaling the tile
                # in place between T.copy and T.gemm) changes what is multiplied
                if not isinstance(o.value, T.BufferRegion):
                    _fail(f"{nm}_shared is modified element-wise ({o.value!r}) besides being filled from global memory")
                cs.append(_elementwise_staging(o, nm))
                raw.append(o)
        if len(cs) != 1:
            _fail(f"expected one T.copy into {nm}_shared, found {len(cs)}")
        c = cs[0]
        used.append(raw[0])
        if c.loop is not g.loop:
            _fail(f"T.copy into {nm}_shared is not in the same K loop as T.gemm")
        src = _buf(c.src)
        if not src.is_global:
            _fail(f"{nm}_shared is not filled from a global tensor")
        if src.dtype != tile.dtype:
            _fail(f"T.copy {src.name}->{nm}_shared changes dtype ({src.dtype}->{tile.dtype})")
        if len(src.shape) != 2:
            _fail(f"global operand {src.name} must be 2-D")
        return src, _tile_origin(c.src, {}, f"T.copy {src.name}->{nm}_shared")

    A_g, a_org = staging(A_s, "A")
    B_g, b_org = staging(B_s, "B")
    bvars = list(tr.block_vars)
    split = []  # [(slice block index, K elements per slice)] when the program cuts K over a grid dimension

    def classify(org, dims_are, nm):
        """dims_are: tuple naming each dim of the global operand, e.g. ('M','K').  Returns block var of the M/N dim."""
        blk = None
        for (v, sc, off), role, tile_ext in zip(org, dims_are, (exp_a if nm == "A" else exp_b)):
            if off != 0:
                _fail(f"{nm}: tile origin has a constant offset {off}")
            if role == "K":
                if isinstance(v, dict):  # {k loop: block_K, slice index: K / split_k}
                    others = [x for x in v if x is not g.loop]
                    if v.get(g.loop) != bK or len(others) != 1 or not any(others[0] is bv for bv in bvars):
                        _fail(f"{nm}: split-K index must be slice*(K/split_k) + k*block_K")
                    cand = (others[0], v[others[0]])
                    if split and not (split[0][0] is cand[0] and split[0][1] == cand[1]):
                        _fail("A and B are sliced differently along K")
                    split[:] = [cand]
                elif v is not g.loop or sc != bK:
                    _fail(f"{nm}: K index must be k*block_K (got {v!r}*{sc})")
            else:
                if isinstance(v, dict):
                    _fail(f"{nm}: {role} index mixes two indices")
                if v is None and sc == 0:
                    blk = None  # single tile along this dimension
                elif any(v is bv for bv in bvars) and sc == tile_ext:
                    blk = v
                else:
                    _fail(f"{nm}: {role} index must be block_index*block_{role} (got {v!r}*{sc}, tile {tile_ext})")
        return blk

    a_dims = ("K", "M") if g.transpose_A else ("M", "K")
    b_dims = ("N", "K") if g.transpose_B else ("K", "N")
    m_var = classify(a_org, a_dims, "A")
    n_var = classify(b_org, b_dims, "B")
    M = cv(A_g.shape[1] if g.transpose_A else A_g.shape[0])
    K = cv(A_g.shape[0] if g.transpose_A else A_g.shape[1])
    N = cv(B_g.shape[0] if g.transpose_B else B_g.shape[1])
    K_b = cv(B_g.shape[1] if g.transpose_B else B_g.shape[0])
    if K != K_b:
        _fail(f"K of A ({K}) != K of B ({K_b})")

    # ---- K loop must cover ceildiv(K, bK)
    ext = kloop.extent
    ext_v = cv(ext)
    split_k = 1
    k_per_cta = K
    if split:
        zvar, k_per_cta = split[0]
        split_k = cv(zvar.extent)
        if zvar is m_var or zvar is n_var or k_per_cta * split_k != K or k_per_cta % bK:
            _fail(f"split-K slices must tile K exactly: {split_k} slices of {k_per_cta} for K = {K}, block_K = {bK}")
    if ext_v != -(-k_per_cta // bK):
        _fail(f"K loop runs {ext_v} iterations, expected ceildiv({k_per_cta},{bK}) = {-(-k_per_cta // bK)}")

    # ---- grid must cover the output
    def grid_extent(v):
        return cv(v.extent)

    if m_var is not None and grid_extent(m_var) != -(-M // bM):
        _fail(f"grid dimension of {m_var!r} is {grid_extent(m_var)}, expected ceildiv(M, block_M) = {-(-M // bM)}")
    if n_var is not None and grid_extent(n_var) != -(-N // bN):
        _fail(f"grid dimension of {n_var!r} is {grid_extent(n_var)}, expected ceildiv(N, block_N) = {-(-N // bN)}")
    if m_var is None and M > bM or n_var is None and N > bN:
        _fail("a single tile does not cover the output and no block index walks it")

    # ---- accumulator initialisation: T.clear before the loop, or clear_accum = (k == 0)
    pre = tr.ops[:tr.ops.index(g)]
    clears = [o for o in pre if isinstance(o, T.ClearOp) and o.buffer is C_acc and o.value == 0.0 and o.loop is None]
    cleared = bool(clears)
    used.extend(clears)
    ca = g.clear_accum
    clear_first = False
    if isinstance(ca, BinOp) and ca.op == "==":
        clear_first = (ca.a is g.loop and is_const(ca.b, 0)) or (ca.b is g.loop and is_const(ca.a, 0))
        if not clear_first:
            _fail(f"clear_accum={ca!r} is not `k == 0`")
    elif isinstance(ca, Expr):
        _fail(f"clear_accum={ca!r}")
    elif ca:
        if ext_v != 1:
            _fail("clear_accum=True inside a multi-iteration K loop discards all but the last K block")
        clear_first = True
    if not (cleared or clear_first):
        _fail("the accumulator is never cleared (T.clear(C) before the loop or clear_accum=(k == 0)); "
              "accumulating into an uninitialised tile has no whole-GEMM meaning")

    # ---- epilogue: follow C_acc through on-chip copies to the global store
    post = tr.ops[tr.ops.index(g) + 1:]
    relu = False
    alpha = 1.0
    cur = C_acc
    C_g, c_org = None, None
    accumulate = False
    for o in post:
        if isinstance(o, T.AtomicAddOp):
            # partial tile added into C: element-wise `T.atomic_add(C[by*bM + i, bx*bN + j], C_shared[i, j])`
            # (examples/gemm_splitk/example_tilelang_gemm_splitk.py:39-40) or the tile form atomic_add(C[...], C_shared)
            if o.loop is not None or not isinstance(o.dst, T.BufferRegion) or not o.dst.buffer.is_global:
                _fail("T.atomic_add must add the finished tile into a global tensor after the K loop")
            if relu or alpha != 1.0:
                _fail("element-wise epilogue before T.atomic_add")
            if o.parallel_vars:
                pv = tuple(o.parallel_vars)
                if tuple(v.extent for v in pv) != (bM, bN) or not _is_load(o.value, cur, pv):
                    _fail("T.atomic_add must add `tile[i, j]` of the accumulator for the whole tile")
                starts = []
                for d, st in enumerate(o.dst.starts):
                    lin = as_linear(st)
                    if lin is None or lin[0].pop(pv[d], 0) != 1 or any(v in lin[0] for v in pv) or len(lin[0]) > 1:
                        _fail(f"T.atomic_add destination index {st!r} must be tile_origin + parallel index")
                    base = lin[1]
                    for v, c in lin[0].items():
                        base = v * c + lin[1]
                    starts.append(base)
                region = T.BufferRegion(o.dst.buffer, starts, (bM, bN))
            else:
                if _buf(o.value) is not cur:
                    _fail("T.atomic_add must add the buffer that holds the accumulator")
                region = o.dst
            C_g, c_org = region.buffer, _tile_origin(region, {}, f"T.atomic_add ->{region.buffer.name}")
            accumulate = True
            used.append(o)
            break
        if isinstance(o, T.CopyOp):
            if o.loop is not None:
                _fail("T.copy inside the K loop after T.gemm")
            if _buf(o.src) is not cur:
                _fail(f"epilogue copies a buffer ({_buf(o.src).name}) that does not hold the accumulator")
            dst = _buf(o.dst)
            used.append(o)
            if dst.is_global:
                C_g, c_org = dst, _tile_origin(o.dst, {}, f"T.copy ->{dst.name}")
                break
            if tuple(dst.shape) != (bM, bN):
                _fail(f"epilogue staging buffer {dst.name} has shape {dst.shape}, expected {(bM, bN)}")
            cur = dst
        elif isinstance(o, T.ElemStore):
            if o.loop is not None:
                # inside the K loop it runs once per K block (a geometric rescaling of the running sum), which no
                # whole-problem alpha expresses
                _fail("element-wise update of the accumulator inside the K loop (only a post-loop epilogue is fused)")
            used.append(o)
            if o.buffer is not cur or len(o.index) != 2 or len(o.parallel_vars) != 2 or \
                    any(a is not b for a, b in zip(o.index, o.parallel_vars)):
                _fail("element-wise epilogue must be `C[i, j] = f(C[i, j])` over the whole accumulator tile")
            if tuple(v.extent for v in o.parallel_vars) != (bM, bN):
                _fail("element-wise epilogue does not cover the accumulator tile")
            val = o.value

            def is_self(x):
                return isinstance(x, T.ElemLoad) and x.buffer is cur and tuple(x.index) == tuple(o.index)

            if isinstance(val, Call) and val.fn == "max" and ((is_self(val.args[0]) and is_const(val.args[1], 0)) or
                                                              (is_self(val.args[1]) and is_const(val.args[0], 0))):
                if alpha < 0:
                    _fail("relu after a negative scale")
                relu = True  # examples/quickstart.py:42-43
            elif isinstance(val, BinOp) and val.op == "*" and ((is_self(val.a) and isinstance(val.b, Const)) or
                                                               (is_self(val.b) and isinstance(val.a, Const))):
                if relu:
                    _fail("scale after relu (kernel applies alpha before relu)")
                alpha *= float(val.b.value if is_self(val.a) else val.a.value)
            else:
                _fail(f"element-wise epilogue {val!r} (supported: relu, scale by a constant)")
        elif isinstance(o, T.ClearOp):
            _fail("T.clear/T.fill after T.gemm")
    if C_g is None:
        _fail("the accumulator is never copied to a global tensor")
    # Sweep: the pattern above accounts for the staging copies, the gemm, the accumulator clear and the post-loop
    # epilogue chain.  Any other recorded operation (a second write to an operand tile, work after the store, an op
    # type this recogniser does not know) means the program computes something else.
    for o in tr.ops:
        if not any(o is u for u in used):
            where = "inside the K loop" if getattr(o, "loop", None) is not None else "outside the K loop"
            _fail(f"unexpected {type(o).__name__} {where} that is not part of the copy -> gemm -> epilogue -> store pattern")
    if split and not accumulate:
        _fail("a program that cuts K over a grid dimension must add its partial tiles into C (T.atomic_add)")
    if tuple(cv(x) for x in C_g.shape) != (M, N):
        _fail(f"output tensor {C_g.name} has shape {C_g.shape}, expected {(M, N)}")
    for (v, sc, off), want_v, ext_ in zip(c_org, (m_var, n_var), (bM, bN)):
        if off != 0 or (want_v is None and v is not None) or (want_v is not None and (v is not want_v or sc != ext_)):
            _fail("output tile origin does not match the operand tiles' block indices")
    if len({A_g.param_index, B_g.param_index, C_g.param_index}) != 3:
        _fail("A, B and C must be three distinct kernel parameters")
    extra = [p for p in tr.params if p.param_index not in (A_g.param_index, B_g.param_index, C_g.param_index)]
    if extra:
        _fail(f"extra kernel parameters {[p.name for p in extra]}")
    if A_s.dtype not in ("float16", "bfloat16", "float8_e4m3", "float8_e5m2", "float32", "int8"):
        _fail(f"operand dtype {A_s.dtype}")
    if A_s.dtype == "int8":
        if C_acc.dtype != "int32":
            _fail(f"int8 operands accumulate in int32, not {C_acc.dtype}")
        if C_g.dtype not in ("int32", "float32", "float16", "bfloat16"):
            _fail(f"output dtype {C_g.dtype} for int8 operands")
        if C_g.dtype == "int32" and alpha != 1.0:
            _fail("scaled int32 output")
    else:
        if C_acc.dtype == "float16" and A_s.dtype in ("float16", "float8_e4m3", "float8_e5m2"):
            # The reference lets these operand types accumulate in fp16 (src/op/tcgen5_meta.h:45-58,
            # examples/gemm_fp8/example_tilelang_gemm_fp8_sm100.py:76).  This backend keeps the fp32
            # accumulator and rounds once on the way out: closer to the exact product, same output dtype.
            pass
        elif C_acc.dtype != "float32":
            _fail(f"accumulator dtype {C_acc.dtype}: the tcgen05 path accumulates in fp32 "
                  "(north_star; the reference allows fp16 accumulate for fp16/fp8 operands only, tcgen5_meta.h:45-58)")
        if C_g.dtype not in ("float16", "bfloat16", "float32", "float8_e4m3", "float8_e5m2"):
            _fail(f"output dtype {C_g.dtype}")

    if C_acc.dtype == "float16":
        notes.append("accum_dtype float16 widened to float32 (single rounding to the output dtype)")
    if g.wg_wait not in (0, -1):
        notes.append(f"wg_wait={g.wg_wait} ignored (MMA completion is tracked by the kernel's own mbarriers)")
    return GemmProgram(M=M, N=N, K=K, in_dtype=A_s.dtype, out_dtype=C_g.dtype, accum_dtype=C_acc.dtype,
                       trans_A=g.transpose_A, trans_B=g.transpose_B, a_param=A_g.param_index, b_param=B_g.param_index,
                       c_param=C_g.param_index, block_M=bM, block_N=bN, block_K=bK, num_stages=kloop.num_stages,
                       threads=tr.threads if isinstance(tr.threads, int) else None, relu=relu, alpha=alpha,
                       accumulate=accumulate, split_k=split_k,
                       accumulator="tmem" if C_acc.scope == "shared.tmem" else "fragment", rasterization=tr.swizzle,
                       notes=notes)


# ======================================================================================================
# Second program family: fp8 contraction with fine-grained scales ("2x accumulation"),
# examples/deepseek_deepgemm/example_deepgemm_fp8_2xAcc.py:43-83  ->  tlb200_gemm_blockscaled.
@dataclass
class BlockScaledProgram:
    M: int
    N: int
    K: int
    in_dtype: str
    out_dtype: str
    a_param: int
    b_param: int
    c_param: int
    sa_param: int
    sb_param: int
    block_M: int
    block_N: int
    block_K: int
    num_stages: int
    accum_dtype: str = "float32"
    trans_A: bool = False
    trans_B: bool = True
    rasterization: Optional[dict] = None
    notes: List[str] = field(default_factory=list)

    def summary(self) -> str:
        return (f"block-scaled gemm {self.M}x{self.N}x{self.K} {self.in_dtype}->{self.out_dtype} NT, scales per "
                f"(row, 128 K) x (128 N, 128 K), tile-hint {self.block_M}x{self.block_N}x{self.block_K}")


GROUP = 128  # the reference's group_size (example_deepgemm_fp8_2xAcc.py:29)


def _is_load(x, buf, index):
    return isinstance(x, T.ElemLoad) and x.buffer is buf and len(x.index) == len(index) and \
        all(a is b for a, b in zip(x.index, index))


def _mul_operands(e):
    return (e.a, e.b) if isinstance(e, BinOp) and e.op == "*" else None


def recognize_blockscaled(func: "T.PrimFunc") -> BlockScaledProgram:
    tr = func.trace()
    if not tr.in_kernel:
        _fail("no T.Kernel frame")
    gemms = [o for o in tr.ops if isinstance(o, T.GemmOp)]
    if len(gemms) != 1 or gemms[0].loop is None:
        _fail("a block-scaled program has exactly one T.gemm inside the K loop")
    g = gemms[0]
    kloop = next(li for li in tr.loops if li.var is g.loop)
    A_s, B_s, C_l = _buf(g.A), _buf(g.B), _buf(g.C)
    if g.transpose_A or not g.transpose_B:
        _fail("block-scaled programs are A[M,K] x B[N,K]^T (transpose_B=True)")
    if isinstance(g.clear_accum, Expr) or g.clear_accum:
        _fail("block-scaled programs clear the partial accumulator with T.clear after promoting it")
    if A_s.dtype != B_s.dtype or A_s.dtype not in ("float8_e4m3", "float8_e5m2"):
        _fail(f"block-scaled operands must be fp8 (got {A_s.dtype}, {B_s.dtype})")
    if C_l.scope != "local.fragment" or C_l.dtype != "float32":
        _fail("the partial accumulator must be a float32 fragment")
    (bM, bK), (bN, bK2) = A_s.shape, B_s.shape
    if bK != bK2 or tuple(C_l.shape) != (bM, bN):
        _fail("tile shapes are inconsistent")
    if bK != GROUP:
        _fail(f"block_K must equal the scale group ({GROUP}): every K block is promoted with one scale")
    if bN > GROUP or GROUP % bN:
        _fail(f"block_N must divide {GROUP}: one scales_b entry per tile")

    def staging(tile, nm):
        cs = [o for o in tr.ops if isinstance(o, T.CopyOp) and _buf(o.dst) is tile]
        if len(cs) != 1 or cs[0].loop is not g.loop:
            _fail(f"expected one T.copy into {nm}_shared inside the K loop")
        src = _buf(cs[0].src)
        if not src.is_global or src.dtype != tile.dtype or len(src.shape) != 2:
            _fail(f"{nm}_shared must be filled from a 2-D global tensor of the same dtype")
        return src, _tile_origin(cs[0].src, {}, f"T.copy {src.name}->{nm}_shared")

    A_g, a_org = staging(A_s, "A")
    B_g, b_org = staging(B_s, "B")
    bvars = list(tr.block_vars)

    def origin(org, blk, nm):
        (v0, s0, o0), (v1, s1, o1) = org
        if o0 or o1 or v1 is not g.loop or s1 != bK:
            _fail(f"{nm}: tile origin must be [block*{blk}, k*{bK}]")
        if v0 is None and s0 == 0:
            return None
        if not any(v0 is bv for bv in bvars) or s0 != blk:
            _fail(f"{nm}: row index must be block_index*{blk}")
        return v0

    m_var, n_var = origin(a_org, bM, "A"), origin(b_org, bN, "B")
    (M, K), (N, K_b) = A_g.shape, B_g.shape
    if not all(isinstance(v, int) for v in (M, N, K, K_b)) or K != K_b:
        _fail("static shapes with equal K are required")
    from .language.expr import evaluate
    nkb, nnb = -(-K // GROUP), -(-N // GROUP)
    if int(evaluate(kloop.extent, {})) != nkb:
        _fail(f"K loop must run ceildiv({K},{bK}) = {nkb} iterations")
    for v, ext, blk, nm in ((m_var, M, bM, "M"), (n_var, N, bN, "N")):
        if v is None:
            if ext > blk:
                _fail(f"no block index walks {nm}")
        elif int(evaluate(v.extent, {})) != -(-ext // blk):
            _fail(f"grid does not cover {nm}")

    # ---- the promotion:  S[i] = scales_a[by*bM + i, k] * scales_b[bx*bN // 128, k];  Acc[i,j] += C_l[i,j] * S[i]
    in_loop = [o for o in tr.ops if isinstance(o, T.ElemStore) and o.loop is g.loop]
    acc_st = [o for o in in_loop if len(o.index) == 2]
    if len(acc_st) != 1:
        _fail("expected one element-wise promotion `C_accum[i, j] += C_local[i, j] * scale[i]` in the K loop")
    st = acc_st[0]
    Acc, (pi, pj) = st.buffer, st.parallel_vars if len(st.parallel_vars) == 2 else (None, None)
    if pi is None or st.index[0] is not pi or st.index[1] is not pj or (pi.extent, pj.extent) != (bM, bN) or \
            Acc.scope != "local.fragment" or Acc.dtype != "float32" or tuple(Acc.shape) != (bM, bN):
        _fail("the promotion must cover a float32 fragment of the tile shape")
    v = st.value
    if not (isinstance(v, BinOp) and v.op == "+"):
        _fail("the promotion must be an accumulation (+=)")
    prod = v.b if _is_load(v.a, Acc, (pi, pj)) else v.a if _is_load(v.b, Acc, (pi, pj)) else None
    ops = _mul_operands(prod) if prod is not None else None
    if ops is None:
        _fail("the promotion must add C_local[i, j] * scale[i]")
    part, sc = ops if _is_load(ops[0], C_l, (pi, pj)) else ops[::-1]
    if not _is_load(part, C_l, (pi, pj)) or not (isinstance(sc, T.ElemLoad) and len(sc.index) == 1 and sc.index[0] is pi):
        _fail("the promotion must multiply the T.gemm accumulator by a per-row scale buffer")
    S = sc.buffer
    s_st = [o for o in in_loop if o.buffer is S]
    if len(s_st) != 1 or len(s_st[0].parallel_vars) != 1 or s_st[0].index[0] is not s_st[0].parallel_vars[0] or \
            s_st[0].parallel_vars[0].extent != bM or tuple(S.shape) != (bM,):
        _fail("the per-row scale buffer must be filled once per K block over block_M rows")
    si = s_st[0].parallel_vars[0]
    ops = _mul_operands(s_st[0].value)
    if ops is None or not all(isinstance(x, T.BufferRegion) and x.buffer.is_global for x in ops):
        _fail("the per-row scale must be scales_a[row, k] * scales_b[col_block, k]")
    ra, rb = ops

    def is_row_scale(r):
        lin = as_linear(r.starts[0]) if len(r.starts) == 2 else None
        if lin is None or r.starts[1] is not g.loop:
            return False
        terms, off = lin
        want = {si: 1}
        if m_var is not None:
            want[m_var] = bM
        return off == 0 and terms == want

    if not is_row_scale(ra):
        ra, rb = rb, ra
    if not is_row_scale(ra):
        _fail("scales_a must be indexed [by * block_M + i, k]")
    e = rb.starts[0] if len(rb.starts) == 2 else None
    ok_b = e is not None and rb.starts[1] is g.loop
    if ok_b:
        if isinstance(e, BinOp) and e.op == "//" and is_const(e.b, GROUP):
            lin = as_linear(e.a)
            ok_b = lin is not None and lin[1] == 0 and lin[0] == ({n_var: bN} if n_var is not None else {})
        else:
            lin = as_linear(e)
            ok_b = lin is not None and lin[1] == 0 and (lin[0] == {} if n_var is None else
                                                        (bN == GROUP and lin[0] == {n_var: 1}))
    if not ok_b:
        _fail(f"scales_b must be indexed [bx * block_N // {GROUP}, k]")
    SA, SB = ra.buffer, rb.buffer
    if SA.dtype != "float32" or SB.dtype != "float32" or tuple(SA.shape) != (M, nkb) or tuple(SB.shape) != (nnb, nkb):
        _fail(f"scale tensors must be float32 of shapes {(M, nkb)} and {(nnb, nkb)}")

    # ---- clears: both accumulators before the loop, the partial one again after each promotion
    def cleared(buf, loop):
        return any(isinstance(o, T.ClearOp) and o.buffer is buf and o.value == 0.0 and o.loop is loop for o in tr.ops)

    if not (cleared(C_l, None) and cleared(Acc, None) and cleared(C_l, g.loop)):
        _fail("expected T.clear of both accumulators before the K loop and T.clear(C_local) after each promotion")
    order = [o for o in tr.ops if o is g or o is st or (isinstance(o, T.ClearOp) and o.buffer is C_l and o.loop is g.loop)]
    if order[0] is not g or order[1] is not st:
        _fail("inside the K loop the order must be T.gemm, promotion, T.clear(C_local)")

    # ---- epilogue: Acc -> [C_shared ->] C[by*bM, bx*bN]
    cur, C_g, c_org = Acc, None, None
    for o in tr.ops[tr.ops.index(st) + 1:]:
        if isinstance(o, T.CopyOp) and o.loop is None:
            if _buf(o.src) is not cur:
                _fail("the epilogue must copy the promoted accumulator")
            dst = _buf(o.dst)
            if dst.is_global:
                C_g, c_org = dst, _tile_origin(o.dst, {}, "T.copy ->C")
                break
            cur = dst
        elif isinstance(o, T.ElemStore) and o.loop is None:
            _fail("element-wise epilogues after the promotion are not part of this family")
    if C_g is None or tuple(C_g.shape) != (M, N) or C_g.dtype not in ("bfloat16", "float32", "float16"):
        _fail("the promoted accumulator must be stored to a bfloat16/float16/float32 C[M, N]")
    for (v0, sc0, off), want_v, blk in zip(c_org, (m_var, n_var), (bM, bN)):
        if off != 0 or (want_v is None and v0 is not None) or (want_v is not None and (v0 is not want_v or sc0 != blk)):
            _fail("output tile origin does not match the operand tiles' block indices")
    ids = [A_g.param_index, B_g.param_index, C_g.param_index, SA.param_index, SB.param_index]
    if len(set(ids)) != 5 or len(tr.params) != 5:
        _fail("the kernel parameters must be exactly A, B, C, scales_a, scales_b")
    return BlockScaledProgram(M=M, N=N, K=K, in_dtype=A_s.dtype, out_dtype=C_g.dtype, a_param=ids[0], b_param=ids[1],
                              c_param=ids[2], sa_param=ids[3], sb_param=ids[4], block_M=bM, block_N=bN, block_K=bK,
                              num_stages=kloop.num_stages, rasterization=tr.swizzle)


def recognize(func: "T.PrimFunc", env: Optional[dict] = None):
    """GEMM family first; a program that is not one may still be the block-scaled form."""
    try:
        return recognize_gemm(func, env)
    except NotAGemmProgram as first:
        try:
            return recognize_blockscaled(func)
        except NotAGemmProgram as second:
            # five parameters (A, B, C, two scale tensors) = the author meant the block-scaled form
            raise (second if len(func.params) == 5 else first) from None
