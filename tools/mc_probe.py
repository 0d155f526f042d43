"""Dev probe: A-multicast plans (cfg_cta_group=4) — correctness vs torch and flush-timed rate vs the other plans."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tilelang_mesh_b200 import gemm as G
torch.manual_seed(0)
flush = torch.empty(256 << 20, dtype=torch.int8, device="cuda")
def timed(fn, n=7):
    fn(); fn()
    ts = []
    for _ in range(n):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return sorted(ts)[len(ts) // 2]
for (m, n, k) in [(256, 128, 64), (512, 512, 512), (300, 200, 136), (512, 2048, 4096), (1024, 1024, 1024), (2048, 512, 8192), (2048, 2048, 2048),
                  (1000, 3000, 5000), (4096, 4096, 4096), (8192, 8192, 8192)]:
    for dt in (torch.bfloat16, torch.float8_e4m3fn):
        a = torch.randint(-3, 4, (m, k), device="cuda").float().to(dt)
        b = torch.randint(-3, 4, (n, k), device="cuda").float().to(dt)
        want = a.float() @ b.float().t()
        line = [f"{(m, n, k)} {str(dt)[6:]:>13s}"]
        for bn in (64, 128, 256):
            if n <= bn:
                continue
            for sk in (1,):
                cfg = dict(cta_group=4, block_n=bn, split_k=sk)
                try:
                    c = G.gemm(a, b, out_dtype=torch.float32, **cfg)
                    ok = torch.equal(c, want)
                    cb = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
                    t = timed(lambda: G.gemm(a, b, cb, **cfg))
                    line.append(f"mc/bn{bn}: {'OK' if ok else 'WRONG'} {2.0 * m * n * k / t / 1e9:6.0f}")
                except RuntimeError as e:
                    line.append(f"mc/bn{bn}: ERR {str(e)[:80]}")
        cb = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
        t = timed(lambda: G.gemm(a, b, cb))
        line.append(f"auto: {2.0 * m * n * k / t / 1e9:6.0f}")
        if dt == torch.bfloat16:
            t = timed(lambda: torch.matmul(a, b.t(), out=cb))
            line.append(f"cublas: {2.0 * m * n * k / t / 1e9:6.0f}")
        print(" | ".join(line), flush=True)
