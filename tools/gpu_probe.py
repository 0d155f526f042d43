"""Bring-up probe for the tcgen05 GEMM: runs ONE named case per process (a trapped kernel kills
the CUDA context) and prints a diagnosis detailed enough to debug layouts remotely.

usage: python tools/gpu_probe.py CASE [CASE...]     (each case spawns its own subprocess)
       python tools/gpu_probe.py --list
"""
import json
import os
import subprocess
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

# name: dict(m,n,k, ab, out, ta, tb, cfg..., data=int|randn, time=bool)
CASES = {}


def case(name, m, n, k, ab="bf16", out="f32", ta=False, tb=True, data="int", timeit=False, **cfg):
    CASES[name] = dict(m=m, n=n, k=k, ab=ab, out=out, ta=ta, tb=tb, data=data, timeit=timeit, cfg=cfg)


# --- 1-CTA bring-up, direct epilogue
case("c1_tile_k64", 128, 256, 64, cta_group=1, epilogue=1)
case("c1_tile_k256", 128, 256, 256, cta_group=1, epilogue=1)
case("c1_tile_k1024", 128, 256, 1024, cta_group=1, epilogue=1)
case("c1_multi", 512, 512, 512, cta_group=1, epilogue=1)
case("c1_bn128", 256, 128, 256, cta_group=1, epilogue=1, block_n=128)
case("c1_bn64", 256, 64, 256, cta_group=1, epilogue=1, block_n=64)
case("c1_bn32", 128, 32, 128, cta_group=1, epilogue=1, block_n=32)
case("c1_tails", 200, 300, 136, cta_group=1, epilogue=1)
# --- TMA-store epilogue
case("c1_tma_f32", 256, 256, 128, cta_group=1, epilogue=2)
case("c1_tma_bf16", 256, 512, 256, out="bf16", cta_group=1, epilogue=2)
case("c1_tma_tails", 200, 304, 136, out="bf16", cta_group=1, epilogue=2)
case("c1_direct_bf16", 256, 512, 256, out="bf16", cta_group=1, epilogue=1)
# --- 2-CTA
case("c2_tile_k64", 256, 256, 64, cta_group=2, epilogue=1)
case("c2_tile_k1024", 256, 256, 1024, cta_group=2, epilogue=1)
case("c2_multi", 1024, 1024, 1024, cta_group=2, epilogue=1)
case("c2_tma_bf16", 1024, 1024, 512, out="bf16", cta_group=2, epilogue=2)
case("c2_tails", 300, 200, 136, out="bf16", cta_group=2, epilogue=2)
case("c2_bn128", 512, 128, 256, cta_group=2, epilogue=1, block_n=128)
# --- fp16 / fp8
case("c1_fp16", 256, 256, 256, ab="f16", out="f16", cta_group=1)
case("c1_fp8", 256, 256, 512, ab="e4m3", out="bf16", cta_group=1)
case("c2_fp8", 512, 512, 512, ab="e4m3", out="bf16", cta_group=2)
case("c2_fp8_out8", 512, 512, 512, ab="e4m3", out="e4m3", cta_group=2, data="small")
case("c1_e5m2", 256, 256, 256, ab="e5m2", out="f32", cta_group=1)
# --- transposed operands
case("c1_nn", 256, 256, 256, ta=False, tb=False, cta_group=1)
case("c1_tn", 256, 256, 256, ta=True, tb=True, cta_group=1)
case("c1_tt", 256, 256, 256, ta=True, tb=False, cta_group=1)
case("c2_nn", 512, 512, 256, ta=False, tb=False, cta_group=2)
case("c2_tt", 512, 512, 256, ta=True, tb=False, cta_group=2)
case("c1_fp8_nn", 256, 256, 256, ab="e4m3", out="f32", ta=False, tb=False, cta_group=1)
# --- random data, big
case("c1_rand_2k", 2048, 2048, 2048, out="bf16", data="randn", cta_group=1)
case("c2_rand_2k", 2048, 2048, 2048, out="bf16", data="randn", cta_group=2)
case("c1_8k", 8192, 8192, 8192, out="bf16", data="randn", timeit=True, cta_group=1)
case("c2_8k", 8192, 8192, 8192, out="bf16", data="randn", timeit=True, cta_group=2)
case("c2_8k_fp8", 8192, 8192, 8192, ab="e4m3", out="bf16", data="randn", timeit=True, cta_group=2)
case("auto_8k", 8192, 8192, 8192, out="bf16", data="randn", timeit=True)
# --- 512-wide tiles (two accumulators per tile, single-buffered TMEM)
case("w512_c2", 1024, 1024, 512, cta_group=2, block_n=512)
case("w512_c2_bf16", 1024, 2048, 512, out="bf16", cta_group=2, block_n=512)
case("w512_c2_tails", 700, 900, 136, out="bf16", cta_group=2, block_n=512)
case("w512_c2_direct", 700, 900, 136, out="bf16", cta_group=2, block_n=512, epilogue=1)
case("w512_c2_e4", 1024, 1024, 512, out="bf16", cta_group=2, block_n=512, epi_warps=4)
case("w512_c1", 512, 1024, 512, cta_group=1, block_n=512)
case("w384_c2", 512, 768, 256, cta_group=2, block_n=384)
case("w512_fp8", 1024, 1024, 1024, ab="e4m3", out="bf16", cta_group=2, block_n=512)
case("w512_nn", 512, 1024, 256, ta=False, tb=False, cta_group=2, block_n=512)
case("w512_tt", 512, 1024, 256, ta=True, tb=False, cta_group=2, block_n=512)
case("w512_rand", 2048, 2048, 2048, out="bf16", data="randn", cta_group=2, block_n=512)
case("w512_8k", 8192, 8192, 8192, out="bf16", data="randn", timeit=True, cta_group=2, block_n=512)
case("w512_8k_e4", 8192, 8192, 8192, out="bf16", data="randn", timeit=True, cta_group=2, block_n=512, epi_warps=4)
case("w512_8k_fp8", 8192, 8192, 8192, ab="e4m3", out="bf16", data="randn", timeit=True, cta_group=2, block_n=512)


def run_case(name):
    import torch
    from tilelang_mesh_b200 import gemm as G

    c = CASES[name]
    tdt = {"bf16": torch.bfloat16, "f16": torch.float16, "e4m3": torch.float8_e4m3fn,
           "e5m2": torch.float8_e5m2, "f32": torch.float32}
    ab, out = tdt[c["ab"]], tdt[c["out"]]
    m, n, k = c["m"], c["n"], c["k"]
    torch.manual_seed(0)
    dev = "cuda"

    def mk(shape):
        if c["data"] == "int":
            return torch.randint(-3, 4, shape, device=dev).float().to(ab)
        if c["data"] == "small":
            return (torch.randint(-2, 3, shape, device=dev).float() * 0.25).to(ab)
        x = torch.randn(shape, device=dev)
        return (x.half().to(ab) if ab in (torch.float8_e4m3fn, torch.float8_e5m2) else x.to(ab))

    a = mk((k, m) if c["ta"] else (m, k))
    b = mk((n, k) if c["tb"] else (k, n))
    af = a.float().t() if c["ta"] else a.float()
    bf = b.float() if c["tb"] else b.float().t()
    ref = af @ bf.t()
    plan = G.plan(m, n, k, ab, out, c["ta"], c["tb"], **c["cfg"])
    cbuf = torch.full((m, n), float("nan"), device=dev, dtype=torch.float32).to(out)
    t0 = time.time()
    G.gemm(a, b, cbuf, trans_a=c["ta"], trans_b=c["tb"], **c["cfg"])
    torch.cuda.synchronize()
    got = cbuf.float()
    refq = ref.to(out).float()
    diff = (got - refq).abs()
    nan = torch.isnan(got).sum().item()
    tol = 1e-2 if c["data"] == "randn" or out != torch.float32 else 0.0
    bad = (diff > tol * (1 + refq.abs())) | torch.isnan(got)
    nbad = bad.sum().item()
    res = dict(case=name, plan={k_: plan[k_] for k_ in ("cta_group", "block_n", "stages", "grid", "epilogue", "smem_bytes")},
               max_abs=float(diff[~torch.isnan(diff)].max().item()) if (~torch.isnan(diff)).any() else None,
               nan=nan, bad=nbad, total=m * n, ok=nbad == 0, secs=round(time.time() - t0, 3))
    if nbad:
        # which (128-row, 32-col) blocks are wrong + a few samples
        idx = bad.nonzero()[:6].tolist()
        res["first_bad"] = [(i, j, got[i, j].item(), refq[i, j].item()) for i, j in idx]
        rb = (m + 127) // 128
        cb = (n + 31) // 32
        blk = torch.zeros(rb, cb, dtype=torch.int32)
        bi = bad.nonzero()
        if len(bi):
            keys = (bi[:, 0] // 128) * cb + (bi[:, 1] // 32)
            uk, cnt = torch.unique(keys, return_counts=True)
            res["bad_blocks_128x32"] = [(int(u) // cb, int(u) % cb, int(cn)) for u, cn in zip(uk[:24], cnt[:24])]
            res["n_bad_blocks"] = int(len(uk))
        # is the result a permutation problem? compare against ref with rows/cols sorted
        res["sum_got"] = float(torch.nan_to_num(got).double().sum().item())
        res["sum_ref"] = float(refq.double().sum().item())
    if c["timeit"] and nbad == 0:
        for _ in range(3):
            G.gemm(a, b, cbuf, trans_a=c["ta"], trans_b=c["tb"], **c["cfg"])
        torch.cuda.synchronize()
        flush = torch.empty(256 * 1024 * 1024, dtype=torch.int8, device=dev)
        ts = []
        for _ in range(10):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            G.gemm(a, b, cbuf, trans_a=c["ta"], trans_b=c["tb"], **c["cfg"])
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
        best, mean = min(ts), sum(ts) / len(ts)
        res["ms_best"], res["ms_mean"] = round(best, 4), round(mean, 4)
        res["tflops_best"] = round(2.0 * m * n * k / best / 1e9, 1)
        res["tflops_mean"] = round(2.0 * m * n * k / mean / 1e9, 1)
        if ab == torch.bfloat16:
            bt = b.t() if c["tb"] else b
            aa = a.t() if c["ta"] else a
            for _ in range(3):
                torch.matmul(aa, bt)
            ts = []
            for _ in range(10):
                flush.zero_()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                torch.matmul(aa, bt)
                e1.record()
                torch.cuda.synchronize()
                ts.append(e0.elapsed_time(e1))
            res["cublas_tflops_best"] = round(2.0 * m * n * k / min(ts) / 1e9, 1)
            res["cublas_tflops_mean"] = round(2.0 * m * n * k / (sum(ts) / len(ts)) / 1e9, 1)
    print("PROBE " + json.dumps(res), flush=True)


if __name__ == "__main__":
    if sys.argv[1:] == ["--list"]:
        print("\n".join(CASES))
    elif len(sys.argv) == 3 and sys.argv[1] == "--one":
        run_case(sys.argv[2])
    else:
        names = sys.argv[1:] or list(CASES)
        for nme in names:
            try:
                p = subprocess.run([sys.executable, os.path.abspath(__file__), "--one", nme], capture_output=True,
                                   text=True, timeout=120)
                out = [l for l in (p.stdout + p.stderr).splitlines() if l.strip()]
                probe = [l for l in out if l.startswith("PROBE ")]
                if probe:
                    print(probe[-1], flush=True)
                else:
                    print(f"FAIL {nme} rc={p.returncode} :: " + " | ".join(out[-6:]), flush=True)
            except subprocess.TimeoutExpired:
                print(f"TIMEOUT {nme}", flush=True)
