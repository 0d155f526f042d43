from .tensor import TensorSupplyType, map_torch_type, torch_assert_close  # noqa: F401
from .target import determine_target  # noqa: F401
