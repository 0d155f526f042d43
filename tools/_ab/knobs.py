import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from tilelang_mesh_b200 import gemm as G
def timed(fn, n=30):
    for _ in range(5): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize(); return e0.elapsed_time(e1) / n
sets = {"bf16": (torch.bfloat16, [dict(), dict(stages=4), dict(epi_warps=4), dict(cta_group=2, block_n=256)]),
        "fp8": (torch.float8_e4m3fn, [dict(), dict(cta_group=2, block_n=512), dict(cta_group=2, block_n=512, stages=4)])}
extra = [eval(x) for x in sys.argv[1:]]
for name, (dt, cfgs) in sets.items():
    a = torch.randn(8192, 8192, device="cuda").to(dt); b = torch.randn(8192, 8192, device="cuda").to(dt); c = torch.empty(8192, 8192, device="cuda", dtype=torch.bfloat16)
    for rep in range(2):
        for cfg in cfgs + extra:
            time.sleep(1.0)
            t = timed(lambda: G.gemm(a, b, c, **cfg))
            print(name, cfg, f"{t*1e3:.1f} us  {2*8192**3/t/1e9:.0f} TFLOP/s", flush=True)
        if dt == torch.bfloat16:
            time.sleep(1.0)
            t = timed(lambda: torch.matmul(a, b.t(), out=c))
            print("cublas", f"{t*1e3:.1f} us  {2*8192**3/t/1e9:.0f} TFLOP/s", flush=True)
