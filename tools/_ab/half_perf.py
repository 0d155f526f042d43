"""Back-to-back (L2-warm) time per call through the C ABI for explicit plans; cuBLAS beside it."""
import ctypes, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from tilelang_mesh_b200 import _abi, gemm as G
lib = _abi.load()
def timed(fn, n=2000):
    for _ in range(50): fn()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): fn()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / n * 1e6
CASES = {
    (512, 2048, 4096): [dict(), dict(cta_group=2, block_n=64), dict(cta_group=3, block_n=128), dict(cta_group=3, block_n=256), dict(cta_group=3, block_n=64, epilogue=1)],
    (1024, 1024, 1024): [dict(), dict(cta_group=3, block_n=128), dict(cta_group=3, block_n=64, epilogue=1), dict(cta_group=2, block_n=64)],
    (2048, 2048, 2048): [dict(), dict(cta_group=3, block_n=256), dict(cta_group=2, block_n=256), dict(cta_group=3, block_n=128)],
    (2048, 512, 8192): [dict(), dict(cta_group=3, block_n=128), dict(cta_group=3, block_n=128, split_k=2), dict(cta_group=3, block_n=256, split_k=2)],
    (256, 4096, 4096): [dict(), dict(cta_group=3, block_n=128), dict(cta_group=3, block_n=256)],
    (128, 4096, 14336): [dict(), dict(cta_group=3, block_n=128, split_k=4), dict(cta_group=3, block_n=128, split_k=2), dict(cta_group=3, block_n=256, split_k=4)],
    (1024, 4096, 4096): [dict(), dict(cta_group=3, block_n=256), dict(cta_group=2, block_n=128)],
    (512, 512, 512): [dict(), dict(cta_group=3, block_n=128), dict(cta_group=3, block_n=64, epilogue=1)],
}
for (m, n, k), cfgs in CASES.items():
    a = torch.randn(m, k, device="cuda").bfloat16(); b = torch.randn(n, k, device="cuda").bfloat16()
    c = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
    stream = torch._C._cuda_getCurrentRawStream(0)
    tl = timed(lambda: torch.matmul(a, b.t(), out=c))
    print(f"== {m}x{n}x{k}: cuBLAS {tl:.2f} us")
    for cfg in cfgs:
        try:
            desc = G.make_desc(m, n, k, G._TORCH2TLB[a.dtype], G._TORCH2TLB[c.dtype], False, True, k, k, n, 1.0, False, **cfg)
            ref = ctypes.byref(desc)
            need = lib.tlb200_gemm_workspace_bytes(ref)
            if need < 0: raise RuntimeError(_abi.last_error() if hasattr(_abi, "last_error") else "rejected")
            ws = torch.empty(max(need, 256), dtype=torch.uint8, device="cuda")
            pa, pb, pc, pw, nw = a.data_ptr(), b.data_ptr(), c.data_ptr(), ws.data_ptr(), ws.numel()
            fn = (lambda: lib.tlb200_gemm_ex(ref, pa, pb, pc, pw, nw, stream)) if need > 0 else (lambda: lib.tlb200_gemm(ref, pa, pb, pc, stream))
            if fn() != 0: raise RuntimeError("launch failed")
            t = timed(fn)
            plan = G.plan_for(a, b, c, **cfg)
            print(f"   {t:7.2f} us  {tl / t:5.2f}x cuBLAS  {cfg}  -> ({plan['cta_group']}, m{plan['block_m']}, n{plan['block_n']}, split {plan['split_k']}, stages {plan['stages']})")
        except Exception as e:
            print(f"   rejected {cfg}: {str(e)[:100]}")
