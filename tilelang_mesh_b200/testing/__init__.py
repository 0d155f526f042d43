"""tilelang.testing helpers used by the reference's GEMM tests (tilelang/testing/__init__.py:30-122)."""
import random
import sys

import numpy as np
import pytest
import torch

from ..utils.tensor import torch_assert_close  # noqa: F401


def set_random_seed(seed: int = 42):
    random.seed(seed)
    np.random.seed(seed)
    torch.manual_seed(seed)
    if torch.cuda.is_available():
        torch.cuda.manual_seed_all(seed)


def _cc():
    if not torch.cuda.is_available():
        return (0, 0)
    return torch.cuda.get_device_capability()


def requires_cuda(fn):
    return pytest.mark.skipif(not torch.cuda.is_available(), reason="CUDA not available")(fn)


def requires_cuda_compute_version(major, minor=0, mode="ge"):
    have, want = _cc(), (major, minor)
    ok = {"ge": have >= want, "gt": have > want, "eq": have == want, "le": have <= want, "lt": have < want}[mode]
    return pytest.mark.skipif(not ok, reason=f"needs compute capability {mode} {major}.{minor}")


def requires_cuda_compute_version_ge(major, minor=0):
    return requires_cuda_compute_version(major, minor, "ge")


def requires_cuda_compute_version_eq(major, minor=0):
    return requires_cuda_compute_version(major, minor, "eq")


def requires_cuda_compute_version_gt(major, minor=0):
    return requires_cuda_compute_version(major, minor, "gt")


def requires_cuda_compute_version_lt(major, minor=0):
    return requires_cuda_compute_version(major, minor, "lt")


def requires_cuda_compute_version_le(major, minor=0):
    return requires_cuda_compute_version(major, minor, "le")


def main():
    frame = sys._getframe(1)
    sys.exit(pytest.main([frame.f_globals["__file__"]] + sys.argv[1:]))
