"""Dev probe: is the K pipeline latency-bound on short problems?  stages sweep + L2 prefetch distance."""
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
for (m, n, k), base in [((512, 2048, 4096), dict(cta_group=1, block_n=64, split_k=1)), ((512, 2048, 4096), dict(cta_group=1, block_n=128, split_k=1)),
                        ((1024, 1024, 1024), dict(cta_group=1, block_n=64, split_k=1)), ((2048, 2048, 2048), dict(cta_group=2, block_n=256, split_k=1)),
                        ((2048, 512, 8192), dict(cta_group=1, block_n=128, split_k=1)), ((8192, 8192, 8192), {}), ((4096, 4096, 4096), {})]:
    a = torch.randn(m, k, device="cuda").bfloat16(); b = torch.randn(n, k, device="cuda").bfloat16()
    c = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
    fl = 2.0 * m * n * k / 1e9
    line = [f"{(m, n, k)} {base}"]
    for st in (0, 3, 5):
        try:
            line.append(f"stages={st}: {fl / timed(lambda: G.gemm(a, b, c, stages=st, **base)):.0f}")
        except RuntimeError as e:
            line.append(f"stages={st}: ERR")
    for pf in (4, 8, 16, 32):
        line.append(f"pf={pf}: {fl / timed(lambda: G.gemm(a, b, c, l2_policy=pf << 8, **base)):.0f}")
    line.append(f"cublas: {fl / timed(lambda: torch.matmul(a, b.t(), out=c)):.0f}")
    print(" | ".join(line), flush=True)
