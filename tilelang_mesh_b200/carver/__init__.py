"""`tilelang.carver` import surface (tilelang/carver/__init__.py).

The reference's carver ("roller") enumerates hardware-aware tile configurations for its code
generator.  This backend has one hand-written kernel whose plan is resolved by a cost model inside
libtlb200.so (`tlb200_gemm_plan_query`), so there is nothing to enumerate: the names the reference's
GEMM examples import exist so that those files load unmodified, and asking the roller for hints
raises NotImplementedError pointing at the plan query.
"""
from . import arch, roller, template  # noqa: F401
from .template import MatmulTemplate  # noqa: F401
