"""tilelang/carver/arch: device descriptions.  Only `driver.get_num_sms()` has a use here
(examples/gemm/example_gemm_persistent.py:57)."""
from . import driver  # noqa: F401


class _Arch:
    def __init__(self, target=None):
        self.target = target
        raise NotImplementedError("tilelang.carver roller hints are not part of the B200 backend; tile plans come "
                                  "from tilelang_mesh_b200.gemm.plan() / tlb200_gemm_plan_query")


class CUDA(_Arch):
    pass


class CDNA(_Arch):
    pass
