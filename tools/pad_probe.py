"""Does any epilogue write beyond column n of a padded C?  (dev probe)"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tilelang_mesh_b200 import gemm as G

for dt, n, pad in ((torch.bfloat16, 500, 12), (torch.bfloat16, 504, 8), (torch.float16, 1001, 7), (torch.float32, 333, 3),
                   (torch.float32, 334, 6), (torch.bfloat16, 250, 6), (torch.bfloat16, 499, 13)):
    for epi in (1, 2):
        a = torch.randn(300, 256, device="cuda", dtype=torch.bfloat16)
        b = torch.randn(n, 256, device="cuda", dtype=torch.bfloat16)
        c = torch.full((300, n + pad), -1.0, device="cuda", dtype=dt)
        try:
            G.gemm(a, b, c[:, :n], trans_b=True, epilogue=epi)
        except RuntimeError as e:
            print(dt, n, pad, "epi", epi, "rejected:", str(e)[:90])
            continue
        badcols = (c[:, n:] != -1).any(dim=0).nonzero().flatten().tolist()
        ok = torch.allclose(c[:, :n].float(), (a.float() @ b.float().t()), rtol=2e-2, atol=2e-2)
        print(dt, n, pad, "epi", epi, "padding columns overwritten:", [n + x for x in badcols], "result ok:", ok)
a = torch.randint(-4, 5, (333, 1152), device="cuda").float().to(torch.float8_e4m3fn)
b = torch.randint(-4, 5, (500, 1152), device="cuda").float().to(torch.float8_e4m3fn)
sa = torch.ones(333, 9, device="cuda")
sb = torch.ones(4, 9, device="cuda")
for dt, pad in ((torch.bfloat16, 12), (torch.float32, 4)):
    c = torch.full((333, 500 + pad), -1.0, device="cuda", dtype=dt)
    G.gemm_blockscaled(a, b, sa, sb, c[:, :500])
    print("blockscaled", dt, "padding columns overwritten:", [500 + x for x in (c[:, 500:] != -1).any(dim=0).nonzero().flatten().tolist()])
