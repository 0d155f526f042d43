"""Host cost of one call: 3000 back-to-back launches of a tiny contraction (GPU time << host time), wall clock."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "compat"))
import torch
from tilelang_mesh_b200 import gemm as G
a = torch.randn(256, 256, device="cuda").bfloat16(); b = torch.randn(256, 256, device="cuda").bfloat16()
c = torch.empty(256, 256, device="cuda", dtype=torch.bfloat16)
def wall(fn, n=3000):
    for _ in range(50):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        fn()
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    return (t1 - t0) / n * 1e6, (t2 - t0) / n * 1e6
print("G.gemm(a, b, c)           : host %.1f us/call, incl. drain %.1f us/call" % wall(lambda: G.gemm(a, b, c)))
print("G.gemm(a, b) (allocates C): host %.1f us/call, incl. drain %.1f us/call" % wall(lambda: G.gemm(a, b)))
print("torch.matmul(out=c)       : host %.1f us/call, incl. drain %.1f us/call" % wall(lambda: torch.matmul(a, b.t(), out=c)))

sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import tilelang  # noqa: E402
import tile_programs as P  # noqa: E402
kern = tilelang.compile(P.nt_fragment(256, 256, 256, 128, 128, 32, dtype="bfloat16"), out_idx=[2])
print("JITKernel(a, b)           : host %.1f us/call, incl. drain %.1f us/call" % wall(lambda: kern(a, b)))
# the same tiny contraction replayed from a CUDA graph of 100 launches: the per-launch cost without the host path
g = torch.cuda.CUDAGraph()
side = torch.cuda.Stream()
side.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(side):
    G.gemm(a, b, c)
    with torch.cuda.graph(g, stream=side):
        for _ in range(100):
            G.gemm(a, b, c)
torch.cuda.current_stream().wait_stream(side)
h, d = wall(lambda: g.replay(), n=100)
print("graph of 100 x G.gemm     : %.2f us per contraction (replay incl. drain)" % (d / 100))
