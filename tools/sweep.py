"""Tile-plan sweep: flush-timed TFLOP/s per (shape, config) to calibrate the automatic plan."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tilelang_mesh_b200 import gemm as G

shapes = [(512, 512, 512), (1024, 1024, 1024), (2048, 2048, 2048), (1024, 4096, 4096), (128, 4096, 14336),
          (64, 8192, 8192), (256, 256, 16384), (2048, 512, 8192), (512, 2048, 4096), (16384, 14336, 4096), (8192, 8192, 8192)]
cfgs = {"auto": {}, "c2n256": dict(cta_group=2, block_n=256, split_k=1), "c2n512": dict(cta_group=2, block_n=512, split_k=1),
        "c1n256": dict(cta_group=1, block_n=256, split_k=1), "c2n128": dict(cta_group=2, block_n=128, split_k=1),
        "c1n128": dict(cta_group=1, block_n=128, split_k=1), "c1n64": dict(cta_group=1, block_n=64, split_k=1),
        "c1n256s8": dict(cta_group=1, block_n=256, split_k=8), "c1n128s4": dict(cta_group=1, block_n=128, split_k=4),
        "c1n128s8": dict(cta_group=1, block_n=128, split_k=8), "c2n256s4": dict(cta_group=2, block_n=256, split_k=4)}
flush = torch.empty(256 << 20, dtype=torch.int8, device="cuda")
for (m, n, k) in shapes:
    a = torch.randn(m, k, device="cuda").bfloat16(); b = torch.randn(n, k, device="cuda").bfloat16()
    c = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
    row = {}
    for nm, cfg in cfgs.items():
        try:
            for _ in range(3): G.gemm(a, b, c, **cfg)
            ts = []
            for _ in range(8):
                flush.zero_()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(); G.gemm(a, b, c, **cfg); e1.record(); torch.cuda.synchronize()
                ts.append(e0.elapsed_time(e1))
            ts.sort()
            row[nm] = round(2.0 * m * n * k / ts[len(ts) // 2] / 1e9, 1)
        except RuntimeError as e:
            row[nm] = None
    for _ in range(3): torch.matmul(a, b.t(), out=c)
    ts = []
    for _ in range(8):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); torch.matmul(a, b.t(), out=c); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    row["cublas"] = round(2.0 * m * n * k / ts[len(ts) // 2] / 1e9, 1)
    p = G.plan(m, n, k)
    print("SWEEP", (m, n, k), json.dumps(row), "auto-plan:", (p["cta_group"], p["block_n"], p["split_k"]), flush=True)
