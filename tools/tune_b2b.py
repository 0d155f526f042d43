"""Measured plans from back-to-back timing through the C ABI (L2-warm where the operands fit; resolution ~0.05 us, where
CUDA events around one launch resolve 2 us).  For every shape the plan space of plan_cache.plan_space is timed; a plan
that beats the cycle model's by more than 3 % (and by twice the spread of the model's plan timed at both ends) goes into
the table.   python tools/tune_b2b.py [--out tilelang_mesh_b200/plans/sm148.json] [--calls 400]"""
import ctypes, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from tilelang_mesh_b200 import _abi, gemm as G, plan_cache as PC

MS = [16, 64, 128, 256, 512, 1024, 2048]
NK = [(4096, 4096), (14336, 4096), (4096, 14336), (8192, 8192), (6144, 4096), (2048, 2048), (1024, 1024)]
out = sys.argv[sys.argv.index("--out") + 1] if "--out" in sys.argv else None
calls = int(sys.argv[sys.argv.index("--calls") + 1]) if "--calls" in sys.argv else 400
lib = _abi.load()
os.environ["TLB200_PLAN_CACHE"] = "off"
stream = torch._C._cuda_getCurrentRawStream(0)
sm = torch.cuda.get_device_properties(0).multi_processor_count

def timed(fn, n):
    for _ in range(20): fn()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): fn()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / n * 1e6

for dt in (torch.bfloat16, torch.float8_e4m3fn):
    for (n, k) in NK:
        b = torch.randn(n, k, device="cuda").to(dt)
        for m in MS:
            a = torch.randn(m, k, device="cuda").to(dt)
            c = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
            res = []
            cands = list(PC.plan_space(m, n, k, a.element_size())) + [{}]
            for cfg in cands:
                desc = G.make_desc(m, n, k, G._TORCH2TLB[a.dtype], G._TORCH2TLB[c.dtype], False, True, k, k, n, 1.0, False, **cfg)
                ref = ctypes.byref(desc)
                need = lib.tlb200_gemm_workspace_bytes(ref)
                if need < 0: continue
                ws = torch.empty(max(need, 256), dtype=torch.uint8, device="cuda")
                pa, pb, pc, pw, nw = a.data_ptr(), b.data_ptr(), c.data_ptr(), ws.data_ptr(), ws.numel()
                fn = (lambda: lib.tlb200_gemm_ex(ref, pa, pb, pc, pw, nw, stream)) if need > 0 else (lambda: lib.tlb200_gemm(ref, pa, pb, pc, stream))
                if fn() != 0: continue
                res.append((timed(fn, calls), cfg))
            autos = [t for t, cfg in res if not cfg]
            auto_us, noise = min(autos), max(autos) / min(autos) - 1.0
            best_us, best_cfg = min(res, key=lambda r: r[0])
            keep = bool(best_cfg) and best_us < (1.0 - max(0.03, 2.0 * noise)) * auto_us
            print(f"{m}x{n}x{k} {str(dt)[6:]}: auto {auto_us:.2f} us (spread {noise:.3f}); best {best_us:.2f} us {best_cfg or 'automatic'} of {len(res)}"
                  + ("  -> table" if keep else ""), flush=True)
            if keep and out:
                PC.store(sm, PC.key(m, n, k, a.dtype, c.dtype, False, True),
                         {"cfg": best_cfg, "us": round(best_us, 2), "auto_us": round(auto_us, 2), "candidates": len(res),
                          "auto_spread": round(noise, 4), "note": "back-to-back through the C ABI (tools/tune_b2b.py)"}, out)
