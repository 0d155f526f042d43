"""Tile-plan sweep: flush-timed TFLOP/s per (shape, dtype, config) to calibrate the automatic plan.
Every config and cuBLAS are timed round-robin (one flush-timed launch each per round) so clock drift hits all alike."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tilelang_mesh_b200 import gemm as G

shapes = [(1024, 1024, 1024), (2048, 2048, 2048), (1024, 4096, 4096), (128, 4096, 14336), (64, 8192, 8192), (16, 4096, 4096),
          (2048, 512, 8192), (512, 2048, 4096), (1000, 3000, 5000), (4096, 4096, 4096), (16384, 14336, 4096),
          (16384, 4096, 14336), (8192, 8192, 8192)]
cfgs = {"auto": {}, "c2n512": dict(cta_group=2, block_n=512, split_k=1), "c2n256": dict(cta_group=2, block_n=256, split_k=1),
        "c1n256": dict(cta_group=1, block_n=256, split_k=1), "c2n128": dict(cta_group=2, block_n=128, split_k=1),
        "c1n128": dict(cta_group=1, block_n=128, split_k=1), "c1n64": dict(cta_group=1, block_n=64, split_k=1),
        "c2n64": dict(cta_group=2, block_n=64, split_k=1), "c1n32": dict(cta_group=1, block_n=32, split_k=1),
        "c2n256s2": dict(cta_group=2, block_n=256, split_k=2), "c2n256s4": dict(cta_group=2, block_n=256, split_k=4),
        "c1n256s4": dict(cta_group=1, block_n=256, split_k=4), "c1n128s2": dict(cta_group=1, block_n=128, split_k=2),
        "c1n128s4": dict(cta_group=1, block_n=128, split_k=4), "c1n128s8": dict(cta_group=1, block_n=128, split_k=8),
        "c1n64s4": dict(cta_group=1, block_n=64, split_k=4), "mc256": dict(cta_group=4, block_n=256, split_k=1),
        "mc128": dict(cta_group=4, block_n=128, split_k=1), "mc256s2": dict(cta_group=4, block_n=256, split_k=2)}
flush = torch.empty(256 << 20, dtype=torch.int8, device="cuda")
one = torch.ones((), device="cuda")
dts = sys.argv[1:] or ["bf16", "fp8"]
for (m, n, k) in shapes:
    for dt in dts:
        if dt == "bf16":
            a = torch.randn(m, k, device="cuda").bfloat16(); b = torch.randn(n, k, device="cuda").bfloat16()
            ref = lambda: torch.matmul(a, b.t(), out=c)
        else:
            if k % 16 or n % 16:
                continue
            a = torch.randn(m, k, device="cuda").half().to(torch.float8_e4m3fn)
            b = torch.randn(n, k, device="cuda").half().to(torch.float8_e4m3fn)
            ref = lambda: torch._scaled_mm(a, b.t(), scale_a=one, scale_b=one, out_dtype=torch.bfloat16)
        c = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
        fns = {}
        for nm, cfg in cfgs.items():
            try:
                G.gemm(a, b, c, **cfg)
                fns[nm] = (lambda cfg=cfg: G.gemm(a, b, c, **cfg))
            except RuntimeError:
                pass
        fns["cublas"] = ref
        for f in fns.values():
            f(); f()
        ts = {nm: [] for nm in fns}
        for _ in range(7):
            for nm, f in fns.items():
                flush.zero_()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(); f(); e1.record(); torch.cuda.synchronize()
                ts[nm].append(e0.elapsed_time(e1))
        row = {nm: round(2.0 * m * n * k / sorted(v)[len(v) // 2] / 1e9) for nm, v in ts.items()}
        p = G.plan(m, n, k, a.dtype, torch.bfloat16)
        best = max((v, nm) for nm, v in row.items() if nm not in ("cublas", "auto"))
        print("SWEEP", (m, n, k), dt, json.dumps(row), "auto-plan:", (p["cta_group"] * p["multicast"], p["block_n"], p["split_k"]),
              "best:", best[1], f"auto/best={row['auto'] / best[0]:.2f} auto/cublas={row['auto'] / row['cublas']:.2f}", flush=True)
