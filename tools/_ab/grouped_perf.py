import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from tilelang_mesh_b200 import gemm as G
def timed(fn, n=300):
    for _ in range(20): fn()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): fn()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / n * 1e6
CFGS = [dict(), dict(cta_group=1, block_n=128), dict(cta_group=1, block_n=256), dict(cta_group=2, block_n=256), dict(cta_group=3, block_n=128), dict(cta_group=3, block_n=256)]
for rows, k, n in [([128] * 64, 4096, 4096), ([96] * 64, 4096, 4096), ([64] * 128, 4096, 2048), ([256] * 32, 4096, 4096), ([512] * 16, 4096, 4096), ([100, 28, 300, 76, 512, 8, 1000, 24] * 4, 4096, 4096), ([128] * 8, 4096, 14336)]:
    a = torch.randn(sum(rows), k, device="cuda").bfloat16(); b = torch.randn(len(rows), n, k, device="cuda").bfloat16()
    c = torch.empty(sum(rows), n, device="cuda", dtype=torch.bfloat16)
    flops = 2.0 * sum(rows) * n * k
    out = []
    for cfg in CFGS:
        try:
            t = timed(lambda: G.gemm_grouped(a, b, rows, c, trans_b=True, **cfg))
            out.append(f"{cfg.get('cta_group', 'auto')}/{cfg.get('block_n', '')}: {t:6.1f} us {flops / t / 1e6:6.0f} TF")
        except Exception as e:
            out.append(f"{cfg}: rejected")
    print(f"{len(rows)} groups, rows {rows[:4]}..., K {k}, N {n}:\n    " + " | ".join(out), flush=True)
