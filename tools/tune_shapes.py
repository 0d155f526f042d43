"""Measure the plan space for a list of shapes and write the winners into a plan table (tilelang_mesh_b200.plan_cache).
  python tools/tune_shapes.py [--out tilelang_mesh_b200/plans/sm148.json] [--verify]
--verify: after tuning, time automatic / cached / best-explicit again and print cached/best (the acceptance check of
SURVEY 8(f) N3: a second run that only reads the table reaches the best explicit plan)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tilelang_mesh_b200 import gemm as G, plan_cache as PC

SHAPES = [(1024, 1024, 1024), (2048, 2048, 2048), (1024, 4096, 4096), (128, 4096, 14336), (64, 8192, 8192), (16, 4096, 4096),
          (2048, 512, 8192), (512, 2048, 4096), (1000, 3000, 5000), (4096, 4096, 4096), (16384, 14336, 4096), (16384, 4096, 14336),
          (8192, 8192, 8192), (2560, 8192, 8192), (5120, 5120, 8192), (4096, 14336, 4096), (4096, 4096, 16384), (256, 4096, 4096)]
out = sys.argv[sys.argv.index("--out") + 1] if "--out" in sys.argv else None
verify = "--verify" in sys.argv
torch.manual_seed(0)
for (m, n, k) in SHAPES:
    for dt in (torch.bfloat16, torch.float8_e4m3fn):
        a = torch.randn(m, k, device="cuda").to(dt); b = torch.randn(n, k, device="cuda").to(dt)
        c = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
        e = PC.tune(a, b, c, path=out, verbose=True)
        if verify:
            PC.reset(); G._desc_cache.clear()
            if out: os.environ["TLB200_PLAN_CACHE"] = out
            flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
            def timed2(c1, c2):  # the two calls timed alternately
                ts = [[], []]
                for _ in range(2): G.gemm(a, b, c, **c1); G.gemm(a, b, c, **c2)
                for _ in range(11):
                    for i, cfg in enumerate((c1, c2)):
                        sink = flush.view(torch.int64).sum(); e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                        e0.record(); G.gemm(a, b, c, **cfg); e1.record(); torch.cuda.synchronize(); ts[i].append(e0.elapsed_time(e1) * 1e3)
                return sorted(ts[0])[5], sorted(ts[1])[5]
            cached, best = timed2({}, e["cfg"])
            print("VERIFY " + json.dumps(dict(shape=[m, n, k], dtype=str(dt)[6:], cfg=e["cfg"], tuned_us=e["us"], auto_us=e["auto_us"],
                                              cached_us=round(cached, 1), explicit_us=round(best, 1), cached_over_best=round(best / cached, 3))), flush=True)
