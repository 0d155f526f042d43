"""Layout helpers (tilelang/layout/).  Shared-memory layouts are fixed by the kernel (128B swizzle for
operands and the epilogue slabs), so these return opaque tags that T.annotate_layout accepts."""


class Layout:
    def __init__(self, kind, buffer=None):
        self.kind, self.buffer = kind, buffer

    def __repr__(self):
        return f"Layout({self.kind})"


def make_swizzled_layout(buffer, k_major: bool = True, allow_pad: bool = True):
    """tilelang/layout/swizzle.py: 128B XOR swizzle chosen by the contiguous extent."""
    return Layout("swizzle", buffer)


def make_tcgen05mma_swizzled_layout(buffer, continuity=None, k_major: bool = True):
    return Layout("tcgen05_swizzle", buffer)


def make_full_bank_swizzled_layout(buffer):
    return Layout("swizzle128", buffer)
