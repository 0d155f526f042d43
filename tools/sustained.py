"""Sustained (power-capped) throughput of GEMM variants next to cuBLAS, with NVML clock/power samples.
usage: python tools/sustained.py [seconds_per_variant] [dtype]"""
import os, sys, threading, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import pynvml
from tilelang_mesh_b200 import gemm as G

secs = float(sys.argv[1]) if len(sys.argv) > 1 else 2.0
dt = sys.argv[2] if len(sys.argv) > 2 else "bf16"
pynvml.nvmlInit()
h = pynvml.nvmlDeviceGetHandleByIndex(0)
m = n = k = 8192
torch.manual_seed(0)
if dt == "fp8":
    a = torch.randn(m, k, device="cuda").half().to(torch.float8_e4m3fn); b = torch.randn(n, k, device="cuda").half().to(torch.float8_e4m3fn)
else:
    a = torch.randn(m, k, device="cuda").bfloat16(); b = torch.randn(n, k, device="cuda").bfloat16()
c = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)

class S(threading.Thread):
    def __init__(s): super().__init__(daemon=True); s.clk=[]; s.pw=[]; s.stop=False
    def run(s):
        while not s.stop:
            s.clk.append(pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)); s.pw.append(pynvml.nvmlDeviceGetPowerUsage(h)/1000.0); time.sleep(0.01)

def run(name, fn):
    for _ in range(10): fn()
    torch.cuda.synchronize()
    # calibrate
    e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): fn()
    e1.record(); torch.cuda.synchronize()
    per=e0.elapsed_time(e1)/20
    iters=max(50,int(secs*1000/per))
    s=S(); s.start()
    e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    s.stop=True; s.join()
    ms=e0.elapsed_time(e1)/iters
    half=len(s.clk)//2
    clk=sorted(s.clk[half:]); pw=sorted(s.pw[half:])
    print("SUST "+json.dumps(dict(name=name, tflops=round(2.0*m*n*k/ms/1e9,1), ms=round(ms,4), burst_tflops=round(2.0*m*n*k/per/1e9,1), iters=iters,
          sm_mhz_med_2nd_half=clk[len(clk)//2] if clk else None, power_w_med=round(pw[len(pw)//2],1) if pw else None)), flush=True)
    time.sleep(1.5)

variants = [("auto", {}), ("pair256", dict(cta_group=2, block_n=256)), ("pair512", dict(cta_group=2, block_n=512)),
            ("mc256", dict(cta_group=4, block_n=256)), ("auto_again", {})]
if len(sys.argv) > 3 and sys.argv[3] == "knobs":
    base = dict(cta_group=2, block_n=256) if dt == "fp8" else dict(cta_group=2, block_n=512)
    variants = [("auto", {})] + [(f"raster{r}", dict(base, raster=r)) for r in (4, 8, 32)] + \
               [(f"stages{st}", dict(base, stages=st)) for st in ((4, 5) if dt == "fp8" else (2, 3, 3))] + [("auto_again", {})]
extra = [v for v in sys.argv[3:]]
if dt == "bf16":
    run("cublas", lambda: torch.matmul(a, b.t(), out=c))
for nm, cfg in variants:
    run(nm, lambda cfg=cfg: G.gemm(a, b, c, **cfg))
if dt == "bf16":
    run("cublas_again", lambda: torch.matmul(a, b.t(), out=c))
