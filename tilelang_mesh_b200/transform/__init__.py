from .pass_config import PassConfigKey  # noqa: F401
