import ctypes, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from tilelang_mesh_b200 import _abi, gemm as G
os.environ["TLB200_PLAN_CACHE"] = "off"
def timed(fn, n):
    for _ in range(10): fn()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): fn()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / n * 1e6
for dt in (torch.bfloat16, torch.float8_e4m3fn):
    for (m, n, k) in [(1024, 8192, 8192), (2048, 8192, 8192), (2048, 14336, 4096), (2048, 4096, 14336), (1024, 14336, 4096), (4096, 4096, 4096), (2048, 6144, 4096), (8192, 8192, 8192)]:
        a = torch.randn(m, k, device="cuda").to(dt); b = torch.randn(n, k, device="cuda").to(dt)
        c = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
        p = G.plan_for(a, b, c)
        line = f"{m}x{n}x{k} {str(dt)[6:10]}: auto=({p['cta_group']},{p['block_n']},mc{p['multicast']},sk{p['split_k']})"
        res = {}
        for rep in range(2):
            for name, cfg in (("auto", {}), ("c2n256", dict(cta_group=2, block_n=256)), ("c4n256", dict(cta_group=4, block_n=256)), ("c2n512", dict(cta_group=2, block_n=512)), ("tail", dict(split_k=-1))):
                try:
                    t = timed(lambda: G.gemm(a, b, c, **cfg), 100)
                    res[name] = min(res.get(name, 1e9), t)
                except Exception:
                    pass
                time.sleep(0.3)
        print(line + "  " + "  ".join(f"{k2} {v:.1f}" for k2, v in res.items()), flush=True)
