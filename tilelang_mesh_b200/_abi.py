"""ctypes binding of libtlb200.so (include/tlb200.h).

This is the analogue of the reference's ctypes adapter
(tilelang/jit/adapter/ctypes/adapter.py:104-107 loads the generated library and calls
``init()``; :195-199 passes raw ``data_ptr()`` values and the current stream).  There is no
fallback: if the shared library is missing or cannot be loaded, importing callers get a
RuntimeError that says how to build it.
"""
from __future__ import annotations

import ctypes
import os
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("TLB200_LIB") or os.path.join(_HERE, "lib", "libtlb200.so")
ABI_VERSION = 2

# tlb200_dtype
FLOAT16, BFLOAT16, FLOAT8_E4M3, FLOAT8_E5M2, FLOAT32, INT8, INT32 = 0, 1, 2, 3, 4, 5, 6
EPI_NONE, EPI_RELU, EPI_ACCUMULATE = 0, 1, 2
STATIC_SCRATCH_BYTES = 32 << 20  # TLB200_STATIC_SCRATCH_BYTES


class GemmDesc(ctypes.Structure):
    _fields_ = [
        ("abi_version", ctypes.c_int32),
        ("m", ctypes.c_int32),
        ("n", ctypes.c_int32),
        ("k", ctypes.c_int32),
        ("ab_dtype", ctypes.c_int32),
        ("c_dtype", ctypes.c_int32),
        ("trans_a", ctypes.c_int32),
        ("trans_b", ctypes.c_int32),
        ("lda", ctypes.c_int64),
        ("ldb", ctypes.c_int64),
        ("ldc", ctypes.c_int64),
        ("alpha", ctypes.c_float),
        ("epilogue", ctypes.c_uint32),
        ("cfg_cta_group", ctypes.c_int32),
        ("cfg_block_n", ctypes.c_int32),
        ("cfg_stages", ctypes.c_int32),
        ("cfg_raster", ctypes.c_int32),
        ("cfg_epilogue", ctypes.c_int32),
        ("reserved", ctypes.c_int32 * 3),
        ("vec_flags", ctypes.c_uint32),
        ("reserved2", ctypes.c_int32 * 3),
    ]


class EpilogueVectors(ctypes.Structure):
    """tlb200_epilogue_vectors (include/tlb200.h): float32 device pointers."""
    _fields_ = [("row_scale", ctypes.c_void_p), ("col_scale", ctypes.c_void_p), ("bias", ctypes.c_void_p)]


VEC_ROW_SCALE, VEC_COL_SCALE, VEC_BIAS_N, VEC_BIAS_M = 1, 2, 4, 8


class ChainDesc(ctypes.Structure):
    """tlb200_chain_desc (include/tlb200.h)."""
    _fields_ = [("abi_version", ctypes.c_int32), ("m", ctypes.c_int32), ("k", ctypes.c_int32), ("n1", ctypes.c_int32),
                ("n2", ctypes.c_int32), ("ab_dtype", ctypes.c_int32), ("c_dtype", ctypes.c_int32), ("relu_mid", ctypes.c_int32),
                ("lda", ctypes.c_int64), ("ldb1", ctypes.c_int64), ("ldb2", ctypes.c_int64), ("ldc", ctypes.c_int64),
                ("alpha1", ctypes.c_float), ("alpha2", ctypes.c_float), ("cfg_stages", ctypes.c_int32),
                ("reserved", ctypes.c_int32 * 3)]


class GemmPlan(ctypes.Structure):
    _fields_ = [
        ("cta_group", ctypes.c_int32),
        ("block_m", ctypes.c_int32),
        ("block_n", ctypes.c_int32),
        ("block_k", ctypes.c_int32),
        ("stages", ctypes.c_int32),
        ("tiles_m", ctypes.c_int32),
        ("tiles_n", ctypes.c_int32),
        ("grid", ctypes.c_int32),
        ("threads", ctypes.c_int32),
        ("smem_bytes", ctypes.c_int32),
        ("raster", ctypes.c_int32),
        ("epilogue", ctypes.c_int32),
        ("tmem_cols", ctypes.c_int32),
        ("instr_desc", ctypes.c_uint32),
        ("umma_n", ctypes.c_int32),
        ("epi_warps", ctypes.c_int32),
        ("split_k", ctypes.c_int32),
        ("multicast", ctypes.c_int32),
        ("workspace_bytes", ctypes.c_int64),
    ]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_ if not k.startswith("reserved")}


# Every symbol include/tlb200.h declares: (restype, argtypes)
class BlockScaledDesc(ctypes.Structure):
    """tlb200_blockscaled_desc (include/tlb200.h)."""
    _fields_ = [("abi_version", ctypes.c_int32), ("m", ctypes.c_int32), ("n", ctypes.c_int32), ("k", ctypes.c_int32),
                ("ab_dtype", ctypes.c_int32), ("c_dtype", ctypes.c_int32), ("group_n", ctypes.c_int32),
                ("group_k", ctypes.c_int32), ("lda", ctypes.c_int64), ("ldb", ctypes.c_int64), ("ldc", ctypes.c_int64),
                ("sa_stride_m", ctypes.c_int64), ("sa_stride_k", ctypes.c_int64), ("sb_stride_n", ctypes.c_int64),
                ("sb_stride_k", ctypes.c_int64), ("cfg_cta_group", ctypes.c_int32), ("cfg_block_n", ctypes.c_int32),
                ("cfg_stages", ctypes.c_int32), ("cfg_raster", ctypes.c_int32), ("reserved", ctypes.c_int32 * 4)]


SYMBOLS = {
    "tlb200_init": (ctypes.c_int, [ctypes.c_int]),
    "tlb200_gemm": (ctypes.c_int, [ctypes.POINTER(GemmDesc), ctypes.c_void_p, ctypes.c_void_p,
                                   ctypes.c_void_p, ctypes.c_void_p]),
    "tlb200_gemm_workspace_bytes": (ctypes.c_int64, [ctypes.POINTER(GemmDesc)]),
    "tlb200_gemm_ex": (ctypes.c_int, [ctypes.POINTER(GemmDesc), ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                      ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p]),
    "tlb200_gemm_fused": (ctypes.c_int, [ctypes.POINTER(GemmDesc), ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                         ctypes.POINTER(EpilogueVectors), ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p]),
    "tlb200_gemm_grouped": (ctypes.c_int, [ctypes.POINTER(GemmDesc), ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                           ctypes.POINTER(ctypes.c_int32), ctypes.c_int, ctypes.c_void_p]),
    "tlb200_gemm_bcast": (ctypes.c_int, [ctypes.POINTER(GemmDesc), ctypes.c_void_p, ctypes.c_void_p,
                                         ctypes.POINTER(ctypes.c_void_p), ctypes.c_int, ctypes.c_void_p]),
    "tlb200_gemm_blockscaled": (ctypes.c_int, [ctypes.POINTER(BlockScaledDesc), ctypes.c_void_p, ctypes.c_void_p,
                                               ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]),
    "tlb200_gemm_chain": (ctypes.c_int, [ctypes.POINTER(ChainDesc)] + [ctypes.c_void_p] * 5),
    "tlb200_gemm_static_scratch": (ctypes.c_int, [ctypes.POINTER(GemmDesc), ctypes.c_void_p, ctypes.c_void_p,
                                                  ctypes.c_void_p, ctypes.c_void_p]),
    "tlb200_gemm_host_workspace_bytes": (ctypes.c_int64, [ctypes.POINTER(GemmDesc)]),
    "tlb200_gemm_host": (ctypes.c_int, [ctypes.POINTER(GemmDesc), ctypes.c_void_p, ctypes.c_void_p,
                                        ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64,
                                        ctypes.c_void_p]),
    "tlb200_gemm_plan_query": (ctypes.c_int, [ctypes.POINTER(GemmDesc), ctypes.c_int,
                                              ctypes.POINTER(GemmPlan)]),
    "tlb200_gemm_plan_query_call": (ctypes.c_int, [ctypes.POINTER(GemmDesc), ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p,
                                                   ctypes.c_void_p, ctypes.c_int64, ctypes.POINTER(GemmPlan)]),
    "tlb200_is_own_device_memory": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int64]),
    "tlb200_last_error": (ctypes.c_char_p, []),
    "tlb200_abi_version": (ctypes.c_int, []),
    "tlb200_launch_count": (ctypes.c_int64, []),
    "tlb200_instr_desc": (ctypes.c_uint32, [ctypes.c_int] * 9),
    "tlb200_tcgen5_meta": (ctypes.c_int, [ctypes.c_int] * 5 + [ctypes.POINTER(ctypes.c_int32)]),
    "tlb200_smem_desc": (ctypes.c_uint64, [ctypes.c_uint32] + [ctypes.c_int] * 5),
}

_lib = None
_lock = threading.Lock()


def load():
    """Load libtlb200.so once and type every exported symbol."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; "
                f"g.build()'` (nvcc, sm_100a). There is no CPU or PyTorch fallback for T.gemm.")
        lib = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(lib, name)  # AttributeError here = header and library disagree
            fn.restype = res
            fn.argtypes = args
        got = lib.tlb200_abi_version()
        if got != ABI_VERSION:
            raise RuntimeError(f"libtlb200.so speaks ABI {got}, Python binding expects {ABI_VERSION}")
        _lib = lib
        return lib


def last_error() -> str:
    return load().tlb200_last_error().decode("utf-8", "replace")


def check(rc: int, what: str):
    """rc != 0 -> RuntimeError(get_last_error()), as cython_wrapper.pyx:279-281."""
    if rc != 0:
        raise RuntimeError(f"{what}: {last_error()}")
