"""tlb200_gemm_chain (intermediate in tensor memory) against the same product as two separate contractions
(this library twice; cuBLAS twice), CUDA events, back to back."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tilelang_mesh_b200 import gemm as G
def timed(fn, n=30):
    for _ in range(5): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize(); return e0.elapsed_time(e1) / n * 1e3
for (m, k, n1, n2) in [(65536, 4096, 64, 4096), (65536, 4096, 128, 4096), (16384, 4096, 256, 4096), (131072, 128, 128, 128), (16384, 8192, 256, 8192)]:
    a = torch.randn(m, k, device="cuda").bfloat16(); b1 = (torch.randn(n1, k, device="cuda") / 32).bfloat16(); b2 = (torch.randn(n2, n1, device="cuda") / 8).bfloat16()
    d = torch.empty(m, n2, device="cuda", dtype=torch.bfloat16); p = torch.empty(m, n1, device="cuda", dtype=torch.bfloat16)
    t_chain = timed(lambda: G.gemm_chain(a, b1, b2, d))
    t_two = timed(lambda: (G.gemm(a, b1, p), G.gemm(p, b2, d)))
    t_lib = timed(lambda: (torch.matmul(a, b1.t(), out=p), torch.matmul(p, b2.t(), out=d)))
    fl = 2.0 * m * n1 * (k + n2)
    print(f"M={m} K={k} N1={n1} N2={n2}: chain {t_chain:8.1f} us ({fl/t_chain/1e6:6.0f} TFLOP/s) | two tlb200_gemm {t_two:8.1f} us | two cuBLAS {t_lib:8.1f} us", flush=True)
