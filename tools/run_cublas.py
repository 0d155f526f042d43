"""Launch torch.matmul (cuBLAS) a few times on one shape (for ncu).  usage: python tools/run_cublas.py [M N K] [iters]"""
import sys, torch
args = [int(x) for x in sys.argv[1:] if x.isdigit()]
m, n, k = args[:3] if len(args) >= 3 else (8192, 8192, 8192)
iters = args[3] if len(args) >= 4 else (args[0] if len(args) == 1 else 3)
torch.manual_seed(0)
a = torch.randn(m, k, device="cuda").bfloat16(); b = torch.randn(n, k, device="cuda").bfloat16(); c = torch.empty(m, n, device="cuda", dtype=torch.bfloat16)
for _ in range(iters + 3): torch.matmul(a, b.t(), out=c)
torch.cuda.synchronize(); print("ok")
