"""tilelang_mesh_b200 — a B200-native (sm_100a) backend for TileLang's dense-contraction path.

The package exposes the slice of the reference's Python surface that the T.gemm / T.copy hot path
needs — `jit`, `compile`, `JITKernel`, `language` (as `T`), `Profiler`, `autotune`, `PassConfigKey`,
`TensorSupplyType`, `layout`, `testing` — over a C-ABI shim (libtlb200.so) holding hand-written
tcgen05 / TMA kernels.  `compat/tilelang` makes it importable as `tilelang`.
"""
__version__ = "0.1.0"

from . import language  # noqa: F401
from . import layout, testing, transform, tvm, utils  # noqa: F401
from .autotuner import autotune  # noqa: F401
from .jit import JITKernel, compile, jit, par_compile  # noqa: F401,A004
from .profiler import Profiler, do_bench  # noqa: F401
from .transform.pass_config import PassConfigKey  # noqa: F401
from .utils.tensor import TensorSupplyType  # noqa: F401
from . import gemm  # noqa: F401  (module: whole-problem host API over the C ABI — gemm.gemm, gemm.plan)


def disable_cache():
    """tilelang.disable_cache(): there is no per-program kernel cache to disable (AOT library)."""


def enable_cache():
    pass
