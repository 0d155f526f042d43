"""Host-buffer schedule piece by piece (a Python replica of tlb200_gemm_host's 2-D arrival schedule):
H2D panels only / + products / + D2H, wall clock per step, next to the library call itself."""
import os, sys, time, ctypes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from cuda.bindings import runtime as rt
from tilelang_mesh_b200 import gemm as G, _abi
M = N = K = 8192
ha = torch.randn(M, K).bfloat16().pin_memory(); hb = torch.randn(N, K).bfloat16().pin_memory(); hc = torch.empty(M, N, dtype=torch.bfloat16).pin_memory()
da = torch.empty(M, K, dtype=torch.bfloat16, device="cuda"); db = torch.empty(N, K, dtype=torch.bfloat16, device="cuda"); dc = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
s_in, s_out, s_c = torch.cuda.Stream(), torch.cuda.Stream(), torch.cuda.Stream()
H2D, D2H = rt.cudaMemcpyKind.cudaMemcpyHostToDevice, rt.cudaMemcpyKind.cudaMemcpyDeviceToHost
def cut(total):
    out, rem = [], total
    while rem > 0:
        r = min(max((rem // 3 + 127) // 128 * 128, 256), 2048, rem); out.append(r); rem -= r
    return out
MODE = {"flat": False, "nodep": False}
def step(do_gemm, do_d2h, panels_a=None, panels_b=None):
    ra, rb = panels_a or cut(M), panels_b or cut(N)
    x = y = 0; i = j = 0
    while i < len(ra) or j < len(rb):
        if i < len(ra) and (x <= y or j >= len(rb)):
            r = ra[i]; rt.cudaMemcpyAsync(da.data_ptr() + x * K * 2, ha.data_ptr() + x * K * 2, r * K * 2, H2D, s_in.cuda_stream)
            ev = torch.cuda.Event(); ev.record(s_in)
            if y > 0 and do_gemm:
                s_c.wait_event(ev)
                with torch.cuda.stream(s_c): G.gemm(da[x:x + r], db[:y], dc[x:x + r, :y])
                if do_d2h:
                    if not MODE["nodep"]:
                        e2 = torch.cuda.Event(); e2.record(s_c); s_out.wait_event(e2)
                    if MODE["flat"]: rt.cudaMemcpyAsync(hc.data_ptr(), dc.data_ptr(), y * 2 * r, D2H, s_out.cuda_stream)
                    else: rt.cudaMemcpy2DAsync(hc.data_ptr() + x * N * 2, N * 2, dc.data_ptr() + x * N * 2, N * 2, y * 2, r, D2H, s_out.cuda_stream)
            x += r; i += 1
        else:
            r = rb[j]; rt.cudaMemcpyAsync(db.data_ptr() + y * K * 2, hb.data_ptr() + y * K * 2, r * K * 2, H2D, s_in.cuda_stream)
            ev = torch.cuda.Event(); ev.record(s_in)
            if x > 0 and do_gemm:
                s_c.wait_event(ev)
                with torch.cuda.stream(s_c): G.gemm(da[:x], db[y:y + r], dc[:x, y:y + r])
                if do_d2h:
                    if not MODE["nodep"]:
                        e2 = torch.cuda.Event(); e2.record(s_c); s_out.wait_event(e2)
                    if MODE["flat"]: rt.cudaMemcpyAsync(hc.data_ptr(), dc.data_ptr(), r * 2 * x, D2H, s_out.cuda_stream)
                    else: rt.cudaMemcpy2DAsync(hc.data_ptr() + y * 2, N * 2, dc.data_ptr() + y * 2, N * 2, r * 2, x, D2H, s_out.cuda_stream)
            y += r; j += 1
    torch.cuda.synchronize()
def t(fn, reps=5):
    fn(); t0 = time.perf_counter()
    for _ in range(reps): fn()
    return (time.perf_counter() - t0) / reps * 1e3
print("panels", cut(M))
print(f"H2D panels only            : {t(lambda: step(False, False)):.2f} ms")
print(f"H2D + products             : {t(lambda: step(True, False)):.2f} ms")
print(f"H2D + products + D2H       : {t(lambda: step(True, True)):.2f} ms")
eq = [1024] * 8
print(f"same, 8 equal panels each  : {t(lambda: step(True, True, eq, eq)):.2f} ms")
big = [2048] * 3 + [1024, 512, 256, 256]
print(f"same, panels {big}: {t(lambda: step(True, True, big, big)):.2f} ms")
MODE["flat"] = True
print(f"schedule with the same bytes copied out as contiguous blocks: {t(lambda: step(True, True)):.2f} ms")
MODE["nodep"] = True
print(f"... and no dependency of the copies out on the products     : {t(lambda: step(True, True)):.2f} ms")
MODE["flat"] = False
print(f"2-D copies out, no dependency on the products               : {t(lambda: step(True, True)):.2f} ms")
MODE["nodep"] = False
wide = [2048] * 4
print(f"A shrinking, B 4 x 2048 (all D2H rows >= 4 KiB): {t(lambda: step(True, True, None, wide)):.2f} ms")
print(f"A 4 x 2048, B 4 x 2048: {t(lambda: step(True, True, wide, wide)):.2f} ms")
def d2h_only():
    ra, rb = cut(M), cut(N); x = y = 0; i = j = 0
    while i < len(ra) or j < len(rb):
        if i < len(ra) and (x <= y or j >= len(rb)):
            r = ra[i]
            if y > 0: rt.cudaMemcpy2DAsync(hc.data_ptr() + x * N * 2, N * 2, dc.data_ptr() + x * N * 2, N * 2, y * 2, r, D2H, s_out.cuda_stream)
            x += r; i += 1
        else:
            r = rb[j]
            if x > 0: rt.cudaMemcpy2DAsync(hc.data_ptr() + y * 2, N * 2, dc.data_ptr() + y * 2, N * 2, r * 2, x, D2H, s_out.cuda_stream)
            y += r; j += 1
    torch.cuda.synchronize()
print(f"the D2H blocks of the schedule alone: {t(d2h_only):.2f} ms")
def bfirst():
    rt.cudaMemcpyAsync(db.data_ptr(), hb.data_ptr(), N * K * 2, H2D, s_in.cuda_stream)
    x = 0
    for r in cut(M):
        rt.cudaMemcpyAsync(da.data_ptr() + x * K * 2, ha.data_ptr() + x * K * 2, r * K * 2, H2D, s_in.cuda_stream)
        ev = torch.cuda.Event(); ev.record(s_in); s_c.wait_event(ev)
        with torch.cuda.stream(s_c): G.gemm(da[x:x + r], db, dc[x:x + r])
        e2 = torch.cuda.Event(); e2.record(s_c); s_out.wait_event(e2)
        rt.cudaMemcpyAsync(hc.data_ptr() + x * N * 2, dc.data_ptr() + x * N * 2, r * N * 2, D2H, s_out.cuda_stream)
        x += r
    torch.cuda.synchronize()
print(f"B first, then shrinking A panels, contiguous D2H: {t(bfirst):.2f} ms")
lib = _abi.load()
d = G.make_desc(M, N, K, _abi.BFLOAT16, _abi.BFLOAT16, 0, 1, K, K, N)
need = lib.tlb200_gemm_host_workspace_bytes(ctypes.byref(d)); ws = torch.empty(need, dtype=torch.uint8, device="cuda")
f = lambda: lib.tlb200_gemm_host(ctypes.byref(d), ha.data_ptr(), hb.data_ptr(), hc.data_ptr(), ws.data_ptr(), need, torch.cuda.current_stream().cuda_stream)
print(f"tlb200_gemm_host           : {t(f):.2f} ms")
def whole():
    rt.cudaMemcpyAsync(da.data_ptr(), ha.data_ptr(), M * K * 2, H2D, s_in.cuda_stream); rt.cudaMemcpyAsync(db.data_ptr(), hb.data_ptr(), N * K * 2, H2D, s_in.cuda_stream)
    rt.cudaMemcpyAsync(hc.data_ptr(), dc.data_ptr(), M * N * 2, D2H, s_out.cuda_stream); torch.cuda.synchronize()
print(f"plain H2D 256 MiB with a concurrent D2H 128 MiB: {t(whole):.2f} ms")
