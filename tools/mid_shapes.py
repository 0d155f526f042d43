"""Run a few mid-size shapes once per config (for an ncu launch list: kernel-only durations vs cuBLAS)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tilelang_mesh_b200 import gemm as G
flush = torch.empty(256 << 20, dtype=torch.int8, device="cuda")
for (m, n, k) in [(512, 2048, 4096), (2048, 512, 8192), (2048, 2048, 2048), (1024, 1024, 1024)]:
    a = torch.randn(m, k, device="cuda").bfloat16(); b = torch.randn(n, k, device="cuda").bfloat16()
    c = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
    for cfg in ({}, dict(cta_group=1, block_n=128, split_k=1), dict(cta_group=2, block_n=256, split_k=4), dict(cta_group=1, block_n=64, split_k=1)):
        for _ in range(2):
            flush.zero_()
            G.gemm(a, b, c, **cfg)
    for _ in range(2):
        flush.zero_()
        torch.matmul(a, b.t(), out=c)
torch.cuda.synchronize()
print("ok")
