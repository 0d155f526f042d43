"""Small set of launches covering every kernel variant once, bit-exact against torch (a quick smoke set; compute-sanitizer is closed on this pool)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tilelang_mesh_b200 import gemm as G
torch.manual_seed(0)
def ints(*shape, dt=torch.bfloat16):
    return torch.randint(-3, 4, shape, device="cuda").float().to(dt)
ok = True
for cfg in ({}, dict(cta_group=1, block_n=64), dict(cta_group=2, block_n=256), dict(cta_group=2, block_n=512), dict(cta_group=4, block_n=128),
            dict(split_k=2), dict(cta_group=1, block_n=128, epilogue=1), dict(cta_group=2, block_n=192, stages=2)):
    for dt in (torch.bfloat16, torch.float8_e4m3fn):
        a, b = ints(300, 520, dt=dt), ints(392, 520, dt=dt)
        c = G.gemm(a, b, out_dtype=torch.float32, **cfg)
        good = torch.equal(c, a.float() @ b.float().t())
        ok &= good
        print("gemm", cfg, str(dt)[6:], "ok" if good else "WRONG", flush=True)
a, b = ints(256, 384, dt=torch.float8_e4m3fn), ints(300, 384, dt=torch.float8_e4m3fn)
sa = torch.ones(256, 3, device="cuda"); sb = torch.ones(3, 3, device="cuda")
for cfg in (dict(cta_group=1, block_n=256), dict(cta_group=1, block_n=128), dict(cta_group=2, block_n=256)):
    c = G.gemm_blockscaled(a, b, sa, sb, out_dtype=torch.float32, **cfg)
    good = torch.equal(c, a.float() @ b.float().t()); ok &= good
    print("blockscaled", cfg, "ok" if good else "WRONG", flush=True)
a = ints(300, 256); bg = ints(3, 256, 128)
c = G.gemm_grouped(a, bg, [100, 0, 200], out_dtype=torch.float32)
want = torch.cat([a[:100].float() @ bg[0].float(), a[100:].float() @ bg[2].float()])
ok &= torch.equal(c, want); print("grouped", "ok" if torch.equal(c, want) else "WRONG")
torch.cuda.synchronize()
print("ALL OK" if ok else "FAILURES")
