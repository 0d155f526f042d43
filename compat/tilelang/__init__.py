"""Drop-in alias: with `compat/` on sys.path, `import tilelang` / `import tilelang.language as T`
resolve to the B200 backend, so the reference's GEMM programs (examples/quickstart.py, examples/gemm,
examples/gemm_sm100, benchmark/matmul, benchmark/matmul_fp8) run unmodified."""
import importlib
import os
import sys

_ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
if _ROOT not in sys.path:
    sys.path.insert(0, _ROOT)

import tilelang_mesh_b200 as _impl  # noqa: E402

for _sub in ("language", "jit", "profiler", "autotuner", "layout", "testing", "transform", "utils", "utils.tensor",
             "utils.target", "transform.pass_config", "carver", "carver.arch", "carver.arch.driver", "carver.template",
             "carver.roller", "carver.roller.rasterization", "tvm"):
    sys.modules[__name__ + "." + _sub] = importlib.import_module("tilelang_mesh_b200." + _sub)
sys.modules[__name__] = _impl
