import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from tilelang_mesh_b200 import gemm as G, plan_cache as PC
SHAPES = [(1024, 1024, 1024), (2048, 2048, 2048), (1024, 4096, 4096), (128, 4096, 14336), (64, 8192, 8192), (16, 4096, 4096),
          (2048, 512, 8192), (512, 2048, 4096), (256, 4096, 4096), (512, 512, 512), (256, 256, 16384), (1000, 3000, 5000)]
out = "gpurun_out/plans_small.json"
torch.manual_seed(0)
for (m, n, k) in SHAPES:
    for dt in (torch.bfloat16, torch.float8_e4m3fn):
        a = torch.randn(m, k, device="cuda").to(dt); b = torch.randn(n, k, device="cuda").to(dt)
        c = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
        e = PC.tune(a, b, c, path=out, verbose=True, reps=7)
        print("   auto plan:", {k2: v for k2, v in G.plan_for(a, b, c).items() if k2 in ("cta_group", "block_n", "split_k", "stages", "multicast")}, flush=True)
