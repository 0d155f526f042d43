"""Launch the fine-grained-scale fp8 contraction a few times (for ncu / timing).  No verification.
usage: python tools/run_blockscaled.py M N K iters [cta_group=1 block_n=256 stages=0]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tilelang_mesh_b200 import gemm as G

m, n, k, iters = (int(x) for x in sys.argv[1:5])
cfg = {kv.split("=")[0]: int(kv.split("=")[1]) for kv in sys.argv[5:]}
torch.manual_seed(0)
a = torch.randn(m, k, device="cuda").half().to(torch.float8_e4m3fn)
b = torch.randn(n, k, device="cuda").half().to(torch.float8_e4m3fn)
sa = torch.rand(m, -(-k // 128), device="cuda") + 0.5
sb = torch.rand(-(-n // 128), -(-k // 128), device="cuda") + 0.5
c = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
for _ in range(3):
    G.gemm_blockscaled(a, b, sa, sb, c, **cfg)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(iters):
    G.gemm_blockscaled(a, b, sa, sb, c, **cfg)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / iters
print(f"run_blockscaled {m}x{n}x{k} cfg={cfg}: {ms:.4f} ms/iter  {2.0*m*n*k/ms/1e9:.1f} TFLOP/s (back-to-back)")
