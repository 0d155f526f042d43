import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import torch
from tilelang_mesh_b200 import plan_cache as PC
for (m, n, k) in [(256, 1024, 1024), (512, 2048, 4096), (4096, 4096, 4096)]:
    a = torch.randn(m, k, device="cuda").bfloat16(); b = torch.randn(n, k, device="cuda").bfloat16()
    c = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
    e = PC.tune(a, b, c, path="gpurun_out/tune_check.json", verbose=True, reps=5)
    print(json.dumps(e))
