"""PCIe probe 2: strided (2-D) device-to-host copies by row width, and the host-buffer schedule piece by piece."""
import os, sys, time, ctypes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from cuda.bindings import runtime as rt
N = 8192
d_c = torch.empty(N, N, dtype=torch.bfloat16, device="cuda"); h_c = torch.empty(N, N, dtype=torch.bfloat16).pin_memory()
s = torch.cuda.Stream()
def t(fn, reps=5):
    fn(); torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(reps): fn()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / reps * 1e3
K = rt.cudaMemcpyKind.cudaMemcpyDeviceToHost
for cols in (128, 256, 512, 1024, 2048, 4096, 8192):
    def f():
        for c0 in range(0, N, cols):
            rt.cudaMemcpy2DAsync(h_c.data_ptr() + c0 * 2, N * 2, d_c.data_ptr() + c0 * 2, N * 2, cols * 2, N, K, s.cuda_stream)
    ms = t(f)
    print(f"D2H 128 MiB as column blocks of {cols:5d} bf16 ({cols*2:6d} B rows): {ms:.2f} ms  {N*N*2/ms/1e6:.1f} GB/s", flush=True)
def g():
    for r0 in range(0, N, 1024):
        rt.cudaMemcpyAsync(h_c.data_ptr() + r0 * N * 2, d_c.data_ptr() + r0 * N * 2, 1024 * N * 2, K, s.cuda_stream)
print(f"D2H 128 MiB as 8 contiguous row panels: {t(g):.2f} ms")
