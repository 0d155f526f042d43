"""tilelang_mesh_b200 — B200-native (sm_100a) backend for TileLang's dense-contraction path."""
__version__ = "0.1.0"
