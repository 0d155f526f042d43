"""CTA 0's cycles from kernel entry to the end of its epilogue, operands in L2 (warm) and not (cold: 512 MB read first),
for a list of shapes in ONE process (a -DTLB200_TRACE build; TLB200_LIB=tools/_ab/libtlb200_trace.so)."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from tilelang_mesh_b200 import _abi, gemm as G
lib = _abi.load()
lib.tlb200_debug_set_trace.argtypes = [ctypes.c_void_p]
SHAPES = [(512, 2048, 4096), (1024, 1024, 1024), (2048, 2048, 2048), (2048, 512, 8192), (256, 4096, 4096), (16, 4096, 4096), (64, 8192, 8192), (512, 512, 512), (1024, 4096, 4096)]
flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
tr = torch.zeros(256 * 16, dtype=torch.int64, device="cuda")
def one(a, b, c, cold):
    tr.zero_()
    if cold: flush.view(torch.int64).sum().item()
    torch.cuda.synchronize()
    lib.tlb200_debug_set_trace(tr.data_ptr())
    G.gemm(a, b, c)
    torch.cuda.synchronize()
    lib.tlb200_debug_set_trace(None)
    r = tr.cpu().view(-1, 16)[0].tolist()
    return r[15] - r[13], r[1] - r[13]   # whole CTA, first K block landed
print(f"{'shape':>20} | warm total / first data | cold total / first data   (cycles, median of 5)")
for (m, n, k) in SHAPES:
    a = torch.randn(m, k, device="cuda").bfloat16(); b = torch.randn(n, k, device="cuda").bfloat16()
    c = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
    for _ in range(3): G.gemm(a, b, c)
    w = sorted(one(a, b, c, False) for _ in range(5))[2]
    cd = sorted(one(a, b, c, True) for _ in range(5))[2]
    print(f"{str((m, n, k)):>20} | {w[0]:7d} / {w[1]:5d} | {cd[0]:7d} / {cd[1]:5d}")
