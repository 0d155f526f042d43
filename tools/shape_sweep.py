"""Automatic plan vs cuBLAS over a grid of shapes (bf16 and fp8), flush-timed medians + back-to-back rate.
Output lines go to profiles/shape_sweep_rN.txt."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tilelang_mesh_b200 import gemm as G

shapes = [(512, 2048, 4096), (2048, 512, 8192), (256, 256, 256), (512, 512, 512), (1024, 1024, 1024), (2048, 2048, 2048), (4096, 4096, 4096), (8192, 8192, 8192),
          (16384, 14336, 4096), (16384, 4096, 14336), (1024, 4096, 4096), (128, 4096, 14336), (64, 8192, 8192),
          (16, 4096, 4096), (256, 256, 16384), (2048, 512, 8192), (4096, 7168, 2048), (8192, 1024, 8192), (1000, 3000, 5000)]
flush = torch.empty(512 << 20, dtype=torch.int8, device="cuda")  # read, not written: no dirty lines left in L2


def med(fn, n=9):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(n):
        _ = flush.view(torch.int64).sum()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


def b2b(fn, n=50):
    for _ in range(3):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


one = torch.ones((), device="cuda")
print(f"{'shape':>22s} {'dtype':>5s} | {'ours':>7s} {'cuBLAS':>7s} {'ratio':>5s} | {'ours b2b':>8s} {'cuBLAS b2b':>10s} | plan (cta_group, block_n, split_k)")
for (m, n, k) in shapes:
    for dt in ("bf16", "fp8"):
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
        mine = lambda: G.gemm(a, b, c)
        fl = 2.0 * m * n * k / 1e9
        try:
            t_ref, tb_ref = med(ref), b2b(ref)
        except Exception as e:  # noqa: BLE001
            t_ref = tb_ref = float("nan")
        t, tb = med(mine), b2b(mine)
        p = G.plan_for(a, b, c)
        print(f"{str((m, n, k)):>22s} {dt:>5s} | {fl / t:7.0f} {fl / t_ref:7.0f} {t_ref / t:5.2f} | {fl / tb:8.0f} {fl / tb_ref:10.0f} | "
              f"{(p['cta_group'], p['block_n'], p['split_k'])}", flush=True)
