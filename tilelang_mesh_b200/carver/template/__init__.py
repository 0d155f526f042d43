"""tilelang/carver/template: roller templates.  See the package docstring."""


class MatmulTemplate:
    def __init__(self, *args, **kwargs):
        raise NotImplementedError("MatmulTemplate (roller hints) is not part of the B200 backend: every GEMM program "
                                  "maps to the same tcgen05 kernel and its plan is resolved by tlb200_gemm_plan_query")
