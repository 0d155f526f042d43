import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200, sm_100a); run with -m gpu")


@pytest.fixture(autouse=True)
def _seed():
    # testing/conftest.py:5-21 of the reference seeds every test the same way.
    import random

    import numpy as np
    import torch
    random.seed(0)
    np.random.seed(0)
    torch.manual_seed(0)
    yield


@pytest.fixture(scope="session")
def built_lib():
    """The in-tree libtlb200.so (built by __graft_entry__.build(); rebuilt here if absent)."""
    from tilelang_mesh_b200 import _abi, build
    if not os.path.exists(_abi.LIB_PATH):
        build.build_library()
    return _abi.load()
