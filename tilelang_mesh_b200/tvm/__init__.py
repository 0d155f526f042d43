"""`from tilelang import tvm` — only the shape-symbol constructor GEMM programs use
(`tvm.te.var("m")`, examples/dynamic_shape/example_dynamic.py:19-21).  There is no TVM behind this backend."""
from ..language import dynamic as _dynamic


class te:  # noqa: N801 - reference spelling tvm.te.var
    @staticmethod
    def var(name, dtype="int32"):
        return _dynamic(name, dtype)


class tir:  # noqa: N801
    Var = staticmethod(lambda name, dtype="int32": _dynamic(name, dtype))
