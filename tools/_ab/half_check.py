import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from tilelang_mesh_b200 import gemm as G
torch.manual_seed(0)
ok = True
for (m, n, k) in [(128, 128, 64), (128, 256, 128), (512, 2048, 4096), (1000, 1200, 520), (1024, 1024, 1024), (300, 200, 136), (65, 128, 64)]:
    for dt in (torch.bfloat16, torch.float8_e4m3fn):
        a = torch.randint(-3, 4, (m, k), device="cuda").float().to(dt); b = torch.randint(-3, 4, (n, k), device="cuda").float().to(dt)
        ref = (a.float() @ b.float().t())
        for bn in (64, 128, 256):
            for extra in ({}, dict(epilogue=1), dict(epi_warps=8)):
                try:
                    c = G.gemm(a, b, out_dtype=torch.float32, cta_group=3, block_n=bn, **extra)
                    torch.cuda.synchronize()
                    bad = (c != ref).sum().item()
                    if bad:
                        ok = False
                        idx = (c != ref).nonzero()[:4].tolist()
                        print(f"MISMATCH {m}x{n}x{k} {dt} bn={bn} {extra}: {bad} wrong, first {idx}")
                    else:
                        print(f"ok {m}x{n}x{k} {str(dt)[6:]} bn={bn} {extra}")
                except RuntimeError as e:
                    print(f"rejected {m}x{n}x{k} bn={bn} {extra}: {str(e)[:120]}")
print("ALL OK" if ok else "FAILURES")
