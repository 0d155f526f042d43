"""PCIe floor for the host-buffer path: pinned H2D / D2H rates alone and concurrently, next to tlb200_gemm_host."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
n = 256 << 20
h_in = torch.empty(n, dtype=torch.uint8, pin_memory=True); h_out = torch.empty(n // 2, dtype=torch.uint8, pin_memory=True)
d_in = torch.empty(n, dtype=torch.uint8, device="cuda"); d_out = torch.empty(n // 2, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def t(fn, reps=5):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps): fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps * 1e3
def h2d():
    with torch.cuda.stream(s1): d_in.copy_(h_in, non_blocking=True)
def d2h():
    with torch.cuda.stream(s2): h_out.copy_(d_out, non_blocking=True)
def both():
    h2d(); d2h()
a, b, c = t(h2d), t(d2h), t(both)
print(f"H2D 256 MiB: {a:.2f} ms ({n / a / 1e6:.1f} GB/s) | D2H 128 MiB: {b:.2f} ms ({n / 2 / b / 1e6:.1f} GB/s) | both concurrently: {c:.2f} ms")
print(f"floor for one 8192^3 bf16 step through host buffers (H2D 256 MiB with the D2H hidden): {a:.2f} ms = {2 * 8192**3 / a / 1e9:.0f} TFLOP/s; with both directions sharing: {c:.2f} ms = {2 * 8192**3 / c / 1e9:.0f} TFLOP/s")
