"""Dev probe: kernel time vs K at fixed (M, N): intercept = fixed cost of a launch, slope = time per K block."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tilelang_mesh_b200 import gemm as G
flush = torch.empty(256 << 20, dtype=torch.int8, device="cuda")
def timed(fn, n=11):
    fn(); fn()
    ts = []
    for _ in range(n):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return sorted(ts)[len(ts) // 2] * 1e3
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
x = torch.zeros(1, device="cuda")
print("empty-ish launch (x.add_): %.1f us" % timed(lambda: x.add_(1)))
cases = [((512, 2048), dict(cta_group=1, block_n=64, split_k=1)), ((512, 2048), dict(cta_group=2, block_n=128, split_k=1)),
         ((2048, 2048), dict(cta_group=2, block_n=256, split_k=1))]
if len(sys.argv) > 1 and sys.argv[1] == "mc":
    cases = [((512, 2048), dict(cta_group=4, block_n=64, split_k=1)), ((512, 2048), dict(cta_group=2, block_n=64, split_k=1)),
             ((512, 2048), dict(cta_group=4, block_n=128, split_k=1)), ((512, 2048), dict(cta_group=1, block_n=128, split_k=1)),
             ((512, 2048), dict(cta_group=1, block_n=32, split_k=1))]
for (m, n), cfg in cases:
    for k in (64, 2048, 4096, 8192):
        a = torch.randn(m, k, device="cuda").bfloat16(); b = torch.randn(n, k, device="cuda").bfloat16()
        c = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
        t = timed(lambda: G.gemm(a, b, c, **cfg)); tc = timed(lambda: torch.matmul(a, b.t(), out=c))
        print(f"{(m, n, k)} {cfg}: ours {t:6.1f} us  cublas {tc:6.1f} us", flush=True)
