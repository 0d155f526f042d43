"""Launch the GEMM a few times on fixed shapes (for ncu / timing).  No verification.
usage: python tools/run_gemm.py M N K ab out iters [key=value cfg...]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tilelang_mesh_b200 import gemm as G

m, n, k = (int(x) for x in sys.argv[1:4])
tdt = {"bf16": torch.bfloat16, "f16": torch.float16, "e4m3": torch.float8_e4m3fn, "f32": torch.float32}
ab, out = tdt[sys.argv[4]], tdt[sys.argv[5]]
iters = int(sys.argv[6])
cfg = {kv.split("=")[0]: int(kv.split("=")[1]) for kv in sys.argv[7:]}
torch.manual_seed(0)
a = torch.randn(m, k, device="cuda").half().to(ab) if ab == torch.float8_e4m3fn else torch.randn(m, k, device="cuda").to(ab)
b = torch.randn(n, k, device="cuda").half().to(ab) if ab == torch.float8_e4m3fn else torch.randn(n, k, device="cuda").to(ab)
c = torch.empty(m, n, device="cuda", dtype=out)
for _ in range(3):
    G.gemm(a, b, c, **cfg)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(iters):
    G.gemm(a, b, c, **cfg)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / iters
print(f"run_gemm {m}x{n}x{k} {sys.argv[4]}->{sys.argv[5]} cfg={cfg}: {ms:.4f} ms/iter  {2.0*m*n*k/ms/1e9:.1f} TFLOP/s (back-to-back, no L2 flush)")
