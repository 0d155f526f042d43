"""Dev probe: tail split (cfg_split_k = -1 forces it, 1 switches it off) — correctness and flush-timed rate."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tilelang_mesh_b200 import gemm as G
flush = torch.empty(256 << 20, dtype=torch.int8, device="cuda")
def timed(fn, n=9):
    fn(); fn()
    ts = []
    for _ in range(n):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return sorted(ts)[len(ts) // 2]
for (m, n, k) in [(4096, 4096, 4096), (4096, 14336, 4096), (4096, 4096, 16384), (3000, 5000, 4096), (5120, 5120, 8192), (16384, 14336, 4096), (2560, 8192, 8192)]:
    for dt in (torch.bfloat16, torch.float8_e4m3fn):
        ai = torch.randint(-3, 4, (m, k), device="cuda").float().to(dt); bi = torch.randint(-3, 4, (n, k), device="cuda").float().to(dt)
        want = ai.float() @ bi.float().t()
        a = torch.randn(m, k, device="cuda").half().to(dt); b = torch.randn(n, k, device="cuda").half().to(dt)
        c = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
        fl = 2.0 * m * n * k / 1e9
        line = [f"{(m, n, k)} {str(dt)[6:10]}"]
        for nm, cfg in (("auto", {}), ("pair256 off", dict(cta_group=2, block_n=256, split_k=1)), ("pair256 tail", dict(cta_group=2, block_n=256, split_k=-1)),
                        ("pair128 tail", dict(cta_group=2, block_n=128, split_k=-1))):
            p = G.plan(m, n, k, dt, torch.bfloat16, **cfg)
            ok = torch.equal(G.gemm(ai, bi, out_dtype=torch.float32, **cfg), want)
            line.append(f"{nm}: {'ok' if ok else 'WRONG'} {fl / timed(lambda: G.gemm(a, b, c, **cfg)):.0f} (ws {p['workspace_bytes'] >> 20} MiB)")
        if dt == torch.bfloat16:
            line.append(f"cublas: {fl / timed(lambda: torch.matmul(a, b.t(), out=c)):.0f}")
        print(" | ".join(line), flush=True)
