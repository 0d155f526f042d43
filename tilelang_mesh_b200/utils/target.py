"""Target resolution (tilelang/utils/target.py:76): this backend has exactly one target."""
from __future__ import annotations

SUPPORTED = ("auto", "cuda")


def determine_target(target="auto", return_object: bool = False):
    t = str(target).split()[0].lower() if target is not None else "auto"
    if t in SUPPORTED or t.startswith("cuda") or t.startswith("nvidia"):
        return "cuda -arch=sm_100a"
    raise NotImplementedError(
        f"target {target!r}: the B200 backend emits sm_100a code only — no multi-backend dispatch "
        "(hip / metal / webgpu / llvm / c targets of the reference are out of scope; graphics-API targets "
        "are unavailable on this box)")
