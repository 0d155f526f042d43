"""Host-side GEMM call: torch tensors -> C ABI (tlb200_gemm).

Mirrors what the reference's kernel adapters do around the generated ``call()``:
argument count / device / dtype / shape / stride checks raised as ValueError
(tilelang/jit/adapter/cython/cython_wrapper.pyx:78-156,259-265), output allocation with
``torch.empty`` on the inputs' device (:192-228), torch's *current* raw stream passed through
(:178-185), non-zero return code -> RuntimeError(get_last_error()) (:279-281).
PyTorch is plumbing only (memory + streams); the arithmetic is the CUDA kernel.
"""
from __future__ import annotations

import ctypes
import os
from typing import Optional

import torch

from . import _abi, plan_cache

_PLAN_HINTS = {"cta_group", "block_n", "stages", "raster", "split_k", "epi_warps", "epilogue"}
_sm_counts = {}
_plan_bypass = False  # plan_cache.tune() sets it while it times explicit plans (the table must not override them)


def _sm_count(dev: int) -> int:
    n = _sm_counts.get(dev)
    if n is None:
        n = _sm_counts[dev] = torch.cuda.get_device_properties(dev).multi_processor_count
    return n

_TORCH2TLB = {
    torch.float16: _abi.FLOAT16,
    torch.bfloat16: _abi.BFLOAT16,
    torch.float8_e4m3fn: _abi.FLOAT8_E4M3,
    torch.float8_e5m2: _abi.FLOAT8_E5M2,
    torch.float32: _abi.FLOAT32,
    torch.int8: _abi.INT8,
    torch.int32: _abi.INT32,
}
# float32 operands are multiplied as TF32 (kind::tf32); int8 operands accumulate in int32 (kind::i8)
_AB_DTYPES = (torch.float16, torch.bfloat16, torch.float8_e4m3fn, torch.float8_e5m2, torch.float32, torch.int8)

_inited_devices = set()
_desc_cache = {}  # problem signature -> (descriptor, byref(descriptor), workspace bytes)
_lib = None


def _load_lib():
    global _lib
    _lib = _abi.load()
    return _lib


_workspaces = {}  # (device index, stream handle) -> torch.uint8 scratch for split-K partial sums


def _workspace(dev: int, stream: int, nbytes: int) -> torch.Tensor:
    """Scratch memory is owned by the caller of the C ABI — here torch, one growing buffer per
    (device, stream) so that calls on different streams never share it."""
    key = (dev, stream)
    ws = _workspaces.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = torch.empty(max(nbytes, 1 << 20), dtype=torch.uint8, device=torch.device("cuda", dev))
        _workspaces[key] = ws
    return ws


def _ensure_init(device_index: int):
    if device_index not in _inited_devices:
        lib = _abi.load()
        _abi.check(lib.tlb200_init(device_index), f"tlb200_init({device_index})")
        _inited_devices.add(device_index)


def _row_pitch(t: torch.Tensor, name: str) -> int:
    if t.dim() != 2:
        raise ValueError(f"{name}: expected a 2-D tensor, got {t.dim()}-D")
    if t.shape[1] > 1 and t.stride(1) != 1:
        raise ValueError(f"{name}: innermost dimension must be contiguous (stride {t.stride(1)})")
    if t.shape[0] > 1:
        return t.stride(0)
    return max(t.shape[1], 1)


def make_desc(m, n, k, ab_dtype, c_dtype, trans_a, trans_b, lda, ldb, ldc, alpha=1.0, relu=False,
              cta_group=0, block_n=0, stages=0, raster=0, epilogue=0, epi_warps=0, split_k=0, l2_policy=0,
              accumulate=False, vec_flags=0, kgroup=0) -> _abi.GemmDesc:
    d = _abi.GemmDesc()
    d.abi_version = _abi.ABI_VERSION
    d.m, d.n, d.k = int(m), int(n), int(k)
    d.ab_dtype, d.c_dtype = int(ab_dtype), int(c_dtype)
    d.trans_a, d.trans_b = int(bool(trans_a)), int(bool(trans_b))
    d.lda, d.ldb, d.ldc = int(lda), int(ldb), int(ldc)
    d.alpha = float(alpha)
    d.epilogue = (_abi.EPI_RELU if relu else _abi.EPI_NONE) | (_abi.EPI_ACCUMULATE if accumulate else 0)
    d.cfg_cta_group, d.cfg_block_n, d.cfg_stages = int(cta_group), int(block_n), int(stages)
    d.cfg_raster, d.cfg_epilogue = int(raster), int(epilogue)
    d.reserved[0] = int(epi_warps)
    d.reserved[1] = int(split_k)
    d.reserved[2] = int(l2_policy)
    d.vec_flags = int(vec_flags)
    d.reserved2[0] = int(kgroup)
    return d


def _vector(v, n, name, device):
    if not isinstance(v, torch.Tensor) or v.dtype != torch.float32 or v.dim() != 1 or v.numel() != n or v.device != device:
        raise ValueError(f"gemm: {name} must be a float32 vector of {n} elements on {device}")
    return v if v.is_contiguous() else v.contiguous()


def gemm(a: torch.Tensor, b: torch.Tensor, c: Optional[torch.Tensor] = None, *, trans_a: bool = False,
         trans_b: bool = True, out_dtype: Optional[torch.dtype] = None, alpha: float = 1.0,
         relu: bool = False, row_scale: Optional[torch.Tensor] = None, col_scale: Optional[torch.Tensor] = None,
         bias: Optional[torch.Tensor] = None, bias_per_row: bool = False, **cfg) -> torch.Tensor:
    """C = alpha * op(A) @ op(B) with fp32 accumulation on the tcgen05 path.

    ``row_scale`` [M], ``col_scale`` [N] and ``bias`` ([N], or [M] with ``bias_per_row``) are float32 vectors fused
    into the epilogue: C = cast(relu?(((alpha * acc) * row_scale[m]) * col_scale[n] + bias)) (tlb200_gemm_fused).

    ``trans_a`` / ``trans_b`` have T.gemm's meaning (tilelang/language/gemm.py:105-116):
    A is [K, M] if trans_a else [M, K]; B is [N, K] if trans_b else [K, N].
    ``accumulate=True`` (in ``cfg``) adds the product to the contents of ``c`` instead of overwriting them
    (the reference's T.atomic_add split-K form, examples/gemm_splitk).
    """
    if cfg.get("accumulate") and c is None:
        raise ValueError("gemm: accumulate=True needs the C tensor to accumulate into")
    if not (isinstance(a, torch.Tensor) and isinstance(b, torch.Tensor)):
        raise ValueError("gemm: A and B must be torch tensors")
    if not a.is_cuda or not b.is_cuda:
        raise ValueError("gemm: A and B must live on a CUDA device (there is no CPU fallback)")
    if a.device != b.device:
        raise ValueError(f"gemm: A on {a.device} but B on {b.device}")
    if a.dtype != b.dtype:
        raise ValueError(f"gemm: A is {a.dtype} but B is {b.dtype}")
    if a.dtype not in _AB_DTYPES:
        raise ValueError(f"gemm: unsupported operand dtype {a.dtype}")
    if a.data_ptr() % 16:  # TMA bases are 16-byte aligned; an odd view offset costs one torch copy
        a = a.clone(memory_format=torch.contiguous_format)
    if b.data_ptr() % 16:
        b = b.clone(memory_format=torch.contiguous_format)
    lda, ldb = _row_pitch(a, "A"), _row_pitch(b, "B")
    m, ka = (a.shape[1], a.shape[0]) if trans_a else (a.shape[0], a.shape[1])
    n, kb = (b.shape[0], b.shape[1]) if trans_b else (b.shape[1], b.shape[0])
    if ka != kb:  # language/gemm.py:78-81
        raise ValueError(f"gemm: T.gemm K shape check failed: K_A = {ka}, K_B = {kb}")
    if c is None:
        c = torch.empty((m, n), dtype=out_dtype or (torch.int32 if a.dtype == torch.int8 else a.dtype), device=a.device)
    else:
        if c.device != a.device:
            raise ValueError(f"gemm: C on {c.device} but A on {a.device}")
        if tuple(c.shape) != (m, n):
            raise ValueError(f"gemm: C has shape {tuple(c.shape)}, expected {(m, n)}")
        if out_dtype is not None and c.dtype != out_dtype:
            raise ValueError(f"gemm: C is {c.dtype}, expected {out_dtype}")
    if c.dtype not in _TORCH2TLB:
        raise ValueError(f"gemm: unsupported output dtype {c.dtype}")
    ldc = _row_pitch(c, "C")
    dev = a.device.index if a.device.index is not None else torch.cuda.current_device()
    if dev not in _inited_devices:
        _ensure_init(dev)
    # Descriptor + workspace size are cached per problem signature: filling ~25 ctypes fields and a plan query cost
    # more host time than a small contraction takes on the GPU.
    vec_flags = 0
    if row_scale is not None or col_scale is not None or bias is not None:
        if row_scale is not None:
            row_scale, vec_flags = _vector(row_scale, m, "row_scale", a.device), vec_flags | _abi.VEC_ROW_SCALE
        if col_scale is not None:
            col_scale, vec_flags = _vector(col_scale, n, "col_scale", a.device), vec_flags | _abi.VEC_COL_SCALE
        if bias is not None:
            bias = _vector(bias, m if bias_per_row else n, "bias", a.device)
            vec_flags |= _abi.VEC_BIAS_M if bias_per_row else _abi.VEC_BIAS_N
    key = (m, n, ka, a.dtype, c.dtype, trans_a, trans_b, lda, ldb, ldc, alpha, relu, vec_flags, tuple(sorted(cfg.items())) if cfg else (),
           _plan_bypass)
    hit = _desc_cache.get(key)
    lib = _lib or _load_lib()
    if hit is None:
        plan_cfg = cfg
        if not _plan_bypass and not (_PLAN_HINTS & cfg.keys()):
            # no hints from the caller: a measured, persisted plan for this problem wins over the cycle model
            tuned = plan_cache.lookup(_sm_count(dev), m, n, ka, a.dtype, c.dtype, trans_a, trans_b)
            if tuned:
                plan_cfg = dict(cfg, **tuned)
        desc = make_desc(m, n, ka, _TORCH2TLB[a.dtype], _TORCH2TLB[c.dtype], trans_a, trans_b, lda, ldb,
                         ldc, alpha, relu, vec_flags=vec_flags, **plan_cfg)
        need = lib.tlb200_gemm_workspace_bytes(ctypes.byref(desc))
        if need < 0:
            _abi.check(-1, "tlb200_gemm_workspace_bytes")
        if len(_desc_cache) > 4096:
            _desc_cache.clear()
        hit = _desc_cache[key] = (desc, ctypes.byref(desc), need)
    desc, desc_ref, need = hit
    cur = torch.cuda.current_device()
    if cur != dev:  # the C ABI launches on the current device: switch only when the tensors live elsewhere
        torch.cuda.set_device(dev)
    try:
        stream = torch._C._cuda_getCurrentRawStream(dev)
        if vec_flags:
            vec = _abi.EpilogueVectors(row_scale.data_ptr() if row_scale is not None else None,
                                       col_scale.data_ptr() if col_scale is not None else None,
                                       bias.data_ptr() if bias is not None else None)
            ws = _workspace(dev, stream, need) if need > 0 else None
            rc = lib.tlb200_gemm_fused(desc_ref, a.data_ptr(), b.data_ptr(), c.data_ptr(), ctypes.byref(vec),
                                       ws.data_ptr() if ws is not None else None, ws.numel() if ws is not None else 0, stream)
        elif need > 0:  # split-K partial sums and/or re-pitched copies of operands TMA cannot address
            ws = _workspace(dev, stream, need)
            rc = lib.tlb200_gemm_ex(desc_ref, a.data_ptr(), b.data_ptr(), c.data_ptr(), ws.data_ptr(), ws.numel(), stream)
        else:
            rc = lib.tlb200_gemm(desc_ref, a.data_ptr(), b.data_ptr(), c.data_ptr(), stream)
    finally:
        if cur != dev:
            torch.cuda.set_device(cur)
    if rc != 0:
        _abi.check(rc, "tlb200_gemm")
    return c


def gemm_grouped(a: torch.Tensor, b: torch.Tensor, group_rows, c: Optional[torch.Tensor] = None, *, trans_b: bool = False,
                 out_dtype: Optional[torch.dtype] = None, alpha: float = 1.0, relu: bool = False, **cfg) -> torch.Tensor:
    """Grouped contraction (examples/grouped_gemm/example_grouped_gemm_fwd.py): rows of ``a`` [sum_rows, K] are
    cut into consecutive groups of ``group_rows[g]`` rows, group g is multiplied by ``b[g]``
    ([K, N], or [N, K] with ``trans_b``)."""
    if a.dim() != 2 or b.dim() != 3 or not (a.is_cuda and b.is_cuda) or a.dtype != b.dtype or a.dtype not in _AB_DTYPES:
        raise ValueError("gemm_grouped: need CUDA tensors A [sum_rows, K] and B [groups, ., .] of the same supported dtype")
    rows = [int(x) for x in group_rows]
    if len(rows) != b.shape[0] or sum(rows) != a.shape[0]:
        raise ValueError("gemm_grouped: group_rows must have one entry per B block and sum to A's rows")
    if not b.is_contiguous():
        raise ValueError("gemm_grouped: B must be contiguous")
    k = a.shape[1]
    n, kb = (b.shape[1], b.shape[2]) if trans_b else (b.shape[2], b.shape[1])
    if kb != k:
        raise ValueError(f"gemm_grouped: T.gemm K shape check failed: K_A = {k}, K_B = {kb}")
    if c is None:
        c = torch.empty((a.shape[0], n), dtype=out_dtype or a.dtype, device=a.device)
    dev = a.device.index if a.device.index is not None else torch.cuda.current_device()
    _ensure_init(dev)
    desc = make_desc(a.shape[0], n, k, _TORCH2TLB[a.dtype], _TORCH2TLB[c.dtype], False, trans_b, _row_pitch(a, "A"),
                     b.shape[2], _row_pitch(c, "C"), alpha, relu, **cfg)
    arr = (ctypes.c_int32 * len(rows))(*rows)
    lib = _abi.load()
    with torch.cuda.device(dev):
        rc = lib.tlb200_gemm_grouped(ctypes.byref(desc), a.data_ptr(), b.data_ptr(), c.data_ptr(), arr, len(rows),
                                     torch.cuda.current_stream(dev).cuda_stream)
    _abi.check(rc, "tlb200_gemm_grouped")
    return c


def gemm_blockscaled(a: torch.Tensor, b: torch.Tensor, scales_a: torch.Tensor, scales_b: torch.Tensor,
                     c: Optional[torch.Tensor] = None, *, out_dtype: torch.dtype = torch.bfloat16, cta_group: int = 0,
                     block_n: int = 0, stages: int = 0, raster: int = 0) -> torch.Tensor:
    """fp8 contraction with fine-grained scales (examples/deepseek_deepgemm/example_deepgemm_fp8_2xAcc.py:12-83):
    ``c[m, n] = sum_kb scales_a[m, kb] * scales_b[n // 128, kb] * (a[m, kb-th 128 of K] . b[n, same])``.
    ``a`` [M, K] and ``b`` [N, K] are fp8 (e4m3 or e5m2), ``scales_a`` [M, ceil(K/128)] and ``scales_b``
    [ceil(N/128), ceil(K/128)] fp32 with any strides."""
    if a.dim() != 2 or b.dim() != 2 or not (a.is_cuda and b.is_cuda) or a.dtype != b.dtype or \
            a.dtype not in (torch.float8_e4m3fn, torch.float8_e5m2):
        raise ValueError("gemm_blockscaled: A [M, K] and B [N, K] must be CUDA fp8 tensors of the same dtype")
    m, k = a.shape
    n, kb_ = b.shape
    if kb_ != k:
        raise ValueError(f"gemm_blockscaled: T.gemm K shape check failed: K_A = {k}, K_B = {kb_}")
    nkb, nnb = -(-k // 128), -(-n // 128)
    for nm, t, shp in (("scales_a", scales_a, (m, nkb)), ("scales_b", scales_b, (nnb, nkb))):
        if not t.is_cuda or t.dtype != torch.float32 or tuple(t.shape) != shp:
            raise ValueError(f"gemm_blockscaled: {nm} must be a CUDA float32 tensor of shape {shp}, got "
                             f"{t.dtype} {tuple(t.shape)}")
    if c is None:
        c = torch.empty((m, n), dtype=out_dtype, device=a.device)
    elif tuple(c.shape) != (m, n) or not c.is_cuda:
        raise ValueError(f"gemm_blockscaled: C must be a CUDA tensor of shape {(m, n)}")
    dev = a.device.index if a.device.index is not None else torch.cuda.current_device()
    _ensure_init(dev)
    d = _abi.BlockScaledDesc()
    d.abi_version = _abi.ABI_VERSION
    d.m, d.n, d.k = m, n, k
    d.ab_dtype, d.c_dtype = _TORCH2TLB[a.dtype], _TORCH2TLB[c.dtype]
    d.group_n = d.group_k = 128
    d.lda, d.ldb, d.ldc = _row_pitch(a, "A"), _row_pitch(b, "B"), _row_pitch(c, "C")
    d.sa_stride_m, d.sa_stride_k = (scales_a.stride(0), scales_a.stride(1)) if scales_a.numel() else (0, 0)
    d.sb_stride_n, d.sb_stride_k = (scales_b.stride(0), scales_b.stride(1)) if scales_b.numel() else (0, 0)
    d.cfg_cta_group, d.cfg_block_n, d.cfg_stages, d.cfg_raster = cta_group, block_n, stages, raster
    lib = _abi.load()
    with torch.cuda.device(dev):
        rc = lib.tlb200_gemm_blockscaled(ctypes.byref(d), a.data_ptr(), b.data_ptr(), c.data_ptr(), scales_a.data_ptr(),
                                         scales_b.data_ptr(), torch.cuda.current_stream(dev).cuda_stream)
    _abi.check(rc, "tlb200_gemm_blockscaled")
    return c


def gemm_chain(a: torch.Tensor, b1: torch.Tensor, b2: torch.Tensor, c: Optional[torch.Tensor] = None, *,
               out_dtype: Optional[torch.dtype] = None, alpha1: float = 1.0, alpha2: float = 1.0, relu_mid: bool = False,
               stages: int = 0) -> torch.Tensor:
    """D = alpha2 * cast(relu?(alpha1 * A @ B1^T)) @ B2^T with the [M, N1] intermediate kept in tensor memory
    (tlb200_gemm_chain: the A operand of the second tcgen05.mma comes from TMEM).  ``a`` [M, K], ``b1`` [N1, K],
    ``b2`` [N2, N1], float16 or bfloat16, N1 in {64, 128, 192, 256}."""
    ts = (a, b1, b2)
    if any(t.dim() != 2 or not t.is_cuda for t in ts) or len({t.dtype for t in ts}) != 1 or a.dtype not in (torch.float16, torch.bfloat16):
        raise ValueError("gemm_chain: A [M, K], B1 [N1, K], B2 [N2, N1] must be CUDA float16 / bfloat16 tensors of one dtype")
    if a.shape[1] != b1.shape[1] or b2.shape[1] != b1.shape[0]:
        raise ValueError(f"gemm_chain: T.gemm K shape check failed: A{tuple(a.shape)} B1{tuple(b1.shape)} B2{tuple(b2.shape)}")
    m, k, n1, n2 = a.shape[0], a.shape[1], b1.shape[0], b2.shape[0]
    if c is None:
        c = torch.empty((m, n2), dtype=out_dtype or a.dtype, device=a.device)
    elif tuple(c.shape) != (m, n2) or not c.is_cuda:
        raise ValueError(f"gemm_chain: C must be a CUDA tensor of shape {(m, n2)}")
    dev = a.device.index if a.device.index is not None else torch.cuda.current_device()
    _ensure_init(dev)
    d = _abi.ChainDesc()
    d.abi_version = _abi.ABI_VERSION
    d.m, d.k, d.n1, d.n2 = m, k, n1, n2
    d.ab_dtype, d.c_dtype, d.relu_mid = _TORCH2TLB[a.dtype], _TORCH2TLB[c.dtype], int(bool(relu_mid))
    d.lda, d.ldb1, d.ldb2, d.ldc = _row_pitch(a, "A"), _row_pitch(b1, "B1"), _row_pitch(b2, "B2"), _row_pitch(c, "C")
    d.alpha1, d.alpha2, d.cfg_stages = float(alpha1), float(alpha2), int(stages)
    lib = _abi.load()
    with torch.cuda.device(dev):
        rc = lib.tlb200_gemm_chain(ctypes.byref(d), a.data_ptr(), b1.data_ptr(), b2.data_ptr(), c.data_ptr(),
                                   torch.cuda.current_stream(dev).cuda_stream)
    _abi.check(rc, "tlb200_gemm_chain")
    return c


def gemm_bcast(a: torch.Tensor, b: torch.Tensor, dst_ptrs, ldc: int, c_dtype: torch.dtype, *, alpha: float = 1.0,
               relu: bool = False, **cfg) -> None:
    """C = A[M,K] @ B[N,K]^T stored into SEVERAL matrices at once (tlb200_gemm_bcast): ``dst_ptrs[0]`` is the
    local destination, the rest are the same window of C in peer GPUs' memory (raw addresses valid in this
    process, e.g. from torch symmetric memory).  Row pitch ``ldc`` (elements) and dtype are shared."""
    if not (a.is_cuda and b.is_cuda) or a.dtype != b.dtype or a.dtype not in _AB_DTYPES:
        raise ValueError("gemm_bcast: A and B must be CUDA tensors of the same supported dtype")
    if a.shape[1] != b.shape[1]:
        raise ValueError(f"gemm_bcast: T.gemm K shape check failed: K_A = {a.shape[1]}, K_B = {b.shape[1]}")
    m, k, n = a.shape[0], a.shape[1], b.shape[0]
    dev = a.device.index if a.device.index is not None else torch.cuda.current_device()
    _ensure_init(dev)
    desc = make_desc(m, n, k, _TORCH2TLB[a.dtype], _TORCH2TLB[c_dtype], False, True, _row_pitch(a, "A"), _row_pitch(b, "B"),
                     ldc, alpha, relu, **cfg)
    arr = (ctypes.c_void_p * len(dst_ptrs))(*[int(p) for p in dst_ptrs])
    lib = _abi.load()
    with torch.cuda.device(dev):
        rc = lib.tlb200_gemm_bcast(ctypes.byref(desc), a.data_ptr(), b.data_ptr(), arr, len(dst_ptrs),
                                   torch.cuda.current_stream(dev).cuda_stream)
    _abi.check(rc, "tlb200_gemm_bcast")


def plan(m, n, k, ab_dtype=torch.bfloat16, c_dtype=torch.bfloat16, trans_a=False, trans_b=True,
         sm_count=148, **cfg) -> dict:
    """Resolve the tile plan for a problem without touching a GPU.  Nothing is known about where A lives, so the
    A-multicast plan (``multicast`` = 2) is reported only for an explicit ``cta_group=4``; ``plan_for`` reports the
    plan a call with given tensors runs."""
    lda = m if trans_a else k
    ldb = k if trans_b else n
    desc = make_desc(m, n, k, _TORCH2TLB[ab_dtype], _TORCH2TLB[c_dtype], trans_a, trans_b, lda, ldb, n,
                     **cfg)
    out = _abi.GemmPlan()
    lib = _abi.load()
    _abi.check(lib.tlb200_gemm_plan_query(ctypes.byref(desc), sm_count, ctypes.byref(out)),
               "tlb200_gemm_plan_query")
    return out.as_dict()


def plan_for(a: torch.Tensor, b: torch.Tensor, c: torch.Tensor, *, trans_a: bool = False, trans_b: bool = True,
             workspace_bytes: Optional[int] = None, **cfg) -> dict:
    """The plan ``gemm(a, b, c, ...)`` would run for exactly these tensors on their device (no launch):
    tlb200_gemm_plan_query_call.  ``workspace_bytes`` defaults to what ``gemm`` itself would provide."""
    m, ka = (a.shape[1], a.shape[0]) if trans_a else (a.shape[0], a.shape[1])
    n = b.shape[0] if trans_b else b.shape[1]
    desc = make_desc(m, n, ka, _TORCH2TLB[a.dtype], _TORCH2TLB[c.dtype], trans_a, trans_b, _row_pitch(a, "A"),
                     _row_pitch(b, "B"), _row_pitch(c, "C"), **cfg)
    lib = _abi.load()
    dev = c.device.index if c.device.index is not None else torch.cuda.current_device()
    _ensure_init(dev)
    out = _abi.GemmPlan()
    with torch.cuda.device(dev):
        need = lib.tlb200_gemm_workspace_bytes(ctypes.byref(desc)) if workspace_bytes is None else int(workspace_bytes)
        ws = _workspace(dev, torch._C._cuda_getCurrentRawStream(dev), need) if need > 0 else None
        rc = lib.tlb200_gemm_plan_query_call(ctypes.byref(desc), a.data_ptr(), b.data_ptr(), c.data_ptr(),
                                             ws.data_ptr() if ws is not None else None, need if ws is not None else 0,
                                             ctypes.byref(out))
    _abi.check(rc, "tlb200_gemm_plan_query_call")
    return out.as_dict()


def is_own_device_memory(ptr: int, nbytes: int) -> bool:
    """tlb200_is_own_device_memory on the current device: the condition under which TMA multicast loads are used."""
    return bool(_abi.load().tlb200_is_own_device_memory(ctypes.c_void_p(int(ptr)), int(nbytes)))
