"""Integer / bit-level restatements of the reference's tcgen05 lowering decisions.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  Every function cites the reference code it
restates; all of them are pinned against tests/golden/tcgen05_lowering.json, which was produced
by the reference lineage's own compiler (tests/golden/generate_golden.py).  The product's
equivalents live in libtlb200.so (tlb200_instr_desc / tlb200_tcgen5_meta / tlb200_smem_desc)
and in the kernel's descriptor arithmetic; tests compare the two bit-for-bit.
"""
from __future__ import annotations

_F16, _BF16, _E4M3, _E5M2, _F32 = "float16", "bfloat16", "float8_e4m3", "float8_e5m2", "float32"
_HALF = (_F16, _BF16)
_F8 = (_E4M3, _E5M2)


def tcgen5_mma_meta(M: int, N: int, K: int, ab_dtype: str, c_dtype: str):
    """GetTCGEN5MMAMeta — src/op/tcgen5_meta.h:21-92.

    Returns (atom_m, atom_n, atom_k, enable_ws, enable_2cta) or None when unsupported."""
    ws_ns = (256, 128, 64)
    if ab_dtype in _HALF and c_dtype == _F32:
        if K % 16:
            return None
        if M % 128 == 0:
            for an in range(256, 15, -16):
                if N % an == 0:
                    return (128, an, 16, False, False)
            return None
        for am in (64, 32):
            if M % am == 0:
                for an in ws_ns:
                    if N % an == 0:
                        return (am, an, 16, True, False)
                return None
        return None
    if ab_dtype in _F8 and c_dtype in (_F32, _F16):
        if K % 32:
            return None
        if M % 128 == 0:
            for an in ws_ns:
                if N % an == 0:
                    return (128, an, 32, True, False)
            for an in range(256, 15, -16):
                if N % an == 0:
                    return (128, an, 32, False, True)
            for an in range(256, 7, -8):
                if N % an == 0:
                    return (128, an, 32, False, False)
            return None
        if M % 64 == 0:
            for an in ws_ns:
                if N % an == 0:
                    return (64, an, 32, True, False)
            for an in range(256, 7, -8):
                if N % an == 0:
                    return (64, an, 32, False, False)
            return None
        if M % 32 == 0:
            for an in ws_ns:
                if N % an == 0:
                    return (32, an, 32, True, False)
            return None
    return None


def tcgen5_instr_desc(atom_m: int, atom_n: int, atom_k: int, ab_dtype: str, c_dtype: str,
                      a_is_k_major: bool = True, b_is_k_major: bool = True, scale_in_a: int = 1,
                      scale_in_b: int = 1) -> int:
    """GetTCGEN5InstrDesc — src/op/tcgen5_meta.h:94-172 (field map :142-171)."""
    assert atom_m % 16 == 0 and atom_n % 8 == 0 and atom_k in (16, 32)
    assert scale_in_a in (1, -1) and scale_in_b in (1, -1)
    fmt = {_F16: 0, _BF16: 1, _E4M3: 0, _E5M2: 1}[ab_dtype]
    c_format = {_F16: 0, _F32: 1}[c_dtype]

    def bits(v, start, width):
        return (v & ((1 << width) - 1)) << start

    d = 0
    d |= bits(c_format, 4, 2)
    d |= bits(fmt, 7, 3)
    d |= bits(fmt, 10, 3)
    d |= bits(1 if scale_in_a == -1 else 0, 13, 1)
    d |= bits(1 if scale_in_b == -1 else 0, 14, 1)
    d |= bits(0 if a_is_k_major else 1, 15, 1)
    d |= bits(0 if b_is_k_major else 1, 16, 1)
    d |= bits(atom_n >> 3, 17, 6)
    d |= bits(atom_m >> 4, 24, 5)
    return d


def tcgen05_smem_desc(smem_addr: int, lbo16: int, sbo16: int, base_offset: int = 0,
                      leading_is_absolute: bool = False, swizzle_mode: int = 2) -> int:
    """initialize_tcgen05_descriptor — src/tl_templates/cuda/common.h:481-496.

    Bit layout (Tcgen05SMemDescriptor bitfield): start>>4 @0 (14 b), LBO @16 (14 b), SBO @32 (14 b),
    version=1 @46, base_offset @49 (3 b), lbo_mode @52, layout_type @61 (3 b)."""
    d = (smem_addr >> 4) & 0x3FFF
    d |= (lbo16 & 0x3FFF) << 16
    d |= (sbo16 & 0x3FFF) << 32
    d |= 1 << 46
    d |= (base_offset & 7) << 49
    d |= (1 if leading_is_absolute else 0) << 52
    d |= (swizzle_mode & 7) << 61
    return d


def swizzle_mode_for(continuous_elems: int, elem_bits: int) -> int:
    """makeGemmABLayoutSm100 — src/layout/gemm_layouts.cc:768-788 — as the tcgen05 swizzle enum
    the v2 emitter passes (tcgen05_macro_generator.py:22-27): 2 = 128 B, 4 = 64 B, 6 = 32 B, 0 = none."""
    vec = 128 // elem_bits
    if continuous_elems % (vec * 8) == 0:
        return 2
    if continuous_elems % (vec * 4) == 0:
        return 4
    if continuous_elems % (vec * 2) == 0:
        return 6
    return 0


def k_major_desc_params(block_k: int, elem_bytes: int):
    """K-major operand, swizzled: LBO/SBO and per-MMA K offsets as the v2 emitter computes them
    (tilelang/intrinsics/tcgen05_macro_generator.py:178-190, 232-233, 312-327).

    Returns (lbo16, sbo16, swizzle_mode, [byte offset of MMA ki for a tile of `rows` rows] as a
    function rows -> list)."""
    swz = swizzle_mode_for(block_k, elem_bytes * 8)
    swz_bytes = {2: 128, 4: 64, 6: 32}[swz]
    atom_elems = swz_bytes // elem_bytes
    micro_k = 32 // elem_bytes
    lbo16 = 16 >> 4
    sbo16 = (8 * swz_bytes) >> 4
    k_atoms = block_k // micro_k
    per_atom = max(atom_elems // micro_k, 1)

    def offsets(rows: int):
        return [((ki % per_atom) * micro_k + (ki // per_atom) * rows * atom_elems) * elem_bytes for ki in
                range(k_atoms)]

    return lbo16, sbo16, swz, offsets


def full_bank_swizzle_index(i: int, j: int, elem_bits: int):
    """makeFullBankSwizzleLayout — src/layout/gemm_layouts.cc:421-438: logical (row i, col j) ->
    (tc, ts, index) with the 16-byte chunk XOR-ed by the row within the 8-row group."""
    vec = 128 // elem_bits
    ts, s = divmod(i, 8)
    tc, c = divmod(j // vec, 8)
    v = j % vec
    return tc, ts, v + ((c ^ s) + s * 8) * vec


def tmem_columns(block_n: int) -> int:
    """LowerSharedTmem — src/transform/lower_shared_tmem.cc:142-149: power of two >= 32, <= 512."""
    c = 32
    while c < block_n:
        c *= 2
    assert c <= 512
    return c
