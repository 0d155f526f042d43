import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from tilelang_mesh_b200 import gemm as G
def timed(fn, n=20):
    for _ in range(3): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize(); return e0.elapsed_time(e1) / n
for (m, n, k) in [(16384, 14336, 4096), (16384, 4096, 14336), (8192, 8192, 8192)]:
    a = torch.randn(m, k, device="cuda").bfloat16(); b = torch.randn(n, k, device="cuda").bfloat16(); c = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
    cfgs = [dict(), dict(cta_group=2, block_n=256, raster=16), dict(cta_group=2, block_n=256, raster=8), dict(cta_group=2, block_n=256, raster=4), dict(cta_group=2, block_n=512, raster=8), dict(cta_group=2, block_n=512, raster=32), dict(cta_group=2, block_n=512, raster=4)]
    for rep in range(2):
        for cfg in cfgs:
            time.sleep(1.0)
            t = timed(lambda: G.gemm(a, b, c, **cfg))
            print((m, n, k), cfg, f"{t*1e3:.1f} us  {2*m*n*k/t/1e9:.0f} TFLOP/s", flush=True)
        time.sleep(1.0)
        t = timed(lambda: torch.matmul(a, b.t(), out=c))
        print((m, n, k), "cublas", f"{t*1e3:.1f} us  {2*m*n*k/t/1e9:.0f} TFLOP/s", flush=True)
