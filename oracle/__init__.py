"""oracle/ — CPU restatement of the reference's dense-contraction path.

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs, never by the product package (tilelang_mesh_b200).

Parity status: integer/bit arithmetic (instruction descriptor, smem descriptor, atom selection,
swizzle, K-advance) is pinned against tests/golden/tcgen05_lowering.json, generated from the
reference lineage's own compiler output; the numeric contraction is pinned bit-exactly against
tests/golden/cpu_c_target_matmul.npz (reference C-target program run on CPU) in
product_mode=1 and by tolerance (north_star: rel. err <= 1e-2 bf16/fp16, <= 5e-2 fp8 vs fp32
torch.matmul) in tensor-core mode, because the reference stores no golden vectors for its GPU
path: bit-level parity of the GPU path is UNPINNED.
"""
