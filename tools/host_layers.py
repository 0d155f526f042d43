"""Host time per call of the contraction, by layer (tools/host_layers.py [M N K]) (the GPU work is a 256^3 product, ~3 us: the queue never fills):
   python G.gemm  |  the C ABI through ctypes with a prepared descriptor  |  torch.matmul"""
import ctypes, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tilelang_mesh_b200 import _abi, gemm as G
shapes = [(256, 256, 256), (512, 2048, 4096)] if len(sys.argv) < 4 else [tuple(int(x) for x in sys.argv[1:4])]
for (m, n, k) in shapes:
    a = torch.randn(m, k, device="cuda").bfloat16(); b = torch.randn(n, k, device="cuda").bfloat16()
    c = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
    lib = _abi.load()
    G.gemm(a, b, c); torch.cuda.synchronize()
    desc = G.make_desc(m, n, k, G._TORCH2TLB[a.dtype], G._TORCH2TLB[c.dtype], False, True, k, k, n, 1.0, False)
    ref = ctypes.byref(desc)
    stream = torch._C._cuda_getCurrentRawStream(0)
    pa, pb, pc = a.data_ptr(), b.data_ptr(), c.data_ptr()
    need = lib.tlb200_gemm_workspace_bytes(ref)
    ws = torch.empty(max(need, 1), dtype=torch.uint8, device="cuda")
    def raw():
        if need > 0: return lib.tlb200_gemm_ex(ref, pa, pb, pc, ws.data_ptr(), ws.numel(), stream)
        return lib.tlb200_gemm(ref, pa, pb, pc, stream)
    N = 3000
    res = {}
    for name, fn in (("G.gemm", lambda: G.gemm(a, b, c)), ("C ABI via ctypes", raw), ("torch.matmul", lambda: torch.matmul(a, b.t(), out=c))):
        for _ in range(50): fn()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(N): fn()
        t1 = time.perf_counter()
        torch.cuda.synchronize()
        t2 = time.perf_counter()
        res[name] = ((t1 - t0) / N * 1e6, (t2 - t0) / N * 1e6)
    print(f"{m}x{n}x{k}: " + "; ".join(f"{k2}: host {v[0]:.2f} us/call (with drain {v[1]:.2f})" for k2, v in res.items()))
