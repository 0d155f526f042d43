"""Time tlb200_gemm_blockscaled (fine-grained-scale fp8) against the plain fp8 GEMM of the same shape.
usage: python tools/bench_blockscaled.py [M N K]..."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from tilelang_mesh_b200 import gemm as G  # noqa: E402


def timed(fn, iters=30, flush=None):
    st, en = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for _ in range(3):
        fn()
    if flush is None:
        st.record()
        for _ in range(iters):
            fn()
        en.record()
        torch.cuda.synchronize()
        return st.elapsed_time(en) / iters
    ts = []
    for _ in range(iters):
        flush.zero_()
        st.record()
        fn()
        en.record()
        torch.cuda.synchronize()
        ts.append(st.elapsed_time(en))
    return sum(ts) / len(ts)


def main():
    shapes = [(8192, 8192, 8192), (1024, 1024, 8192), (4096, 7168, 4096), (4096, 4096, 7168), (16384, 4096, 14336)]
    if len(sys.argv) > 3:
        v = [int(x) for x in sys.argv[1:]]
        shapes = [tuple(v[i:i + 3]) for i in range(0, len(v), 3)]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for (M, N, K) in shapes:
        a = torch.randn(M, K, device="cuda", dtype=torch.float16).to(torch.float8_e4m3fn)
        b = torch.randn(N, K, device="cuda", dtype=torch.float16).to(torch.float8_e4m3fn)
        sa = torch.rand(M, -(-K // 128), device="cuda") + 0.5
        sb = torch.rand(-(-N // 128), -(-K // 128), device="cuda") + 0.5
        c = torch.empty(M, N, device="cuda", dtype=torch.bfloat16)
        fl = 2.0 * M * N * K / 1e9
        t_plain = timed(lambda: G.gemm(a, b, c, trans_b=True), flush=flush)
        line = [f"{M}x{N}x{K}: plain fp8 {fl / t_plain:7.0f}"]
        for cg, bn in ((2, 256), (1, 256), (2, 128), (1, 128)):
            try:
                t = timed(lambda: G.gemm_blockscaled(a, b, sa, sb, c, cta_group=cg, block_n=bn), flush=flush)
                tb = timed(lambda: G.gemm_blockscaled(a, b, sa, sb, c, cta_group=cg, block_n=bn))
                line.append(f"cg{cg}/bn{bn} {fl / t:6.0f} (b2b {fl / tb:6.0f})")
            except RuntimeError as e:
                line.append(f"cg{cg}/bn{bn} ERR {str(e)[:60]}")
        print(" | ".join(line) + "  TFLOP/s", flush=True)


if __name__ == "__main__":
    main()
