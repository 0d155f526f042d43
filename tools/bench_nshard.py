"""N-sharded contraction + all-gather of C at G GPUs (north_star's optional variant of C5):
fused (epilogue stores into every rank's C over NVLink, one kernel) vs GEMM + NCCL all-gather
(M-panel overlapped) vs the local GEMM alone.  Launch with torchrun.  Times are CUDA events, max over ranks."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
from tilelang_mesh_b200.parallel import ShardedGemm
from tilelang_mesh_b200 import gemm as G

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
M = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
N = 8192
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 10
K = int(sys.argv[3]) if len(sys.argv) > 3 else 8192  # a small K makes the stores (the NVLink side) the whole cost
torch.manual_seed(0)
A = torch.randn(M, K, device="cuda").bfloat16()
lo, hi = rank * (N // world), (rank + 1) * (N // world)
torch.manual_seed(1)
B = torch.randn(N, K, device="cuda").bfloat16()[lo:hi].contiguous()

def timed(fn):
    for _ in range(2): fn()
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item()

fcfg = {kv.split("=")[0]: int(kv.split("=")[1]) for kv in os.environ.get("NSHARD_CFG", "").split(",") if kv}
sf = ShardedGemm(M, N, K, mode="N", fused=True, fused_cfg=fcfg)
sn = ShardedGemm(M, N, K, mode="N", panel_rows=8192)
cl = torch.empty(M, hi - lo, device="cuda", dtype=torch.bfloat16)
t_f = timed(lambda: sf.run(A, B))
t_n = timed(lambda: sn.run(A, B))
t_g = timed(lambda: G.gemm(A, B, cl))
# the fused kernel alone (no symmetric-memory barriers around it; results are not read): what the kernel itself costs
_ptrs = [int(p) for p in sf._symm_hdl.buffer_ptrs]; _off = sf.lo * 2
_order = [rank] + [g for g in range(world) if g != rank]
t_k = timed(lambda: G.gemm_bcast(A, B, [_ptrs[g] + _off for g in _order], ldc=N, c_dtype=torch.bfloat16, **fcfg))
t_k1 = timed(lambda: G.gemm_bcast(A, B, [_ptrs[rank] + _off], ldc=N, c_dtype=torch.bfloat16, **fcfg))
t_k256 = timed(lambda: G.gemm(A, B, cl, cta_group=2, block_n=256))
same = torch.equal(sf.run(A, B).view(torch.int16), sn.run(A, B).view(torch.int16))
if rank == 0:
    flop = 2.0 * M * N * K
    recv = M * (N - (hi - lo)) * 2
    print("NSHARD " + json.dumps(dict(cfg=fcfg, world=world, M=M, N=N, K=K, ms_fused=round(t_f, 3), ms_gemm_then_nccl_allgather=round(t_n, 3),
          ms_local_gemm_only=round(t_g, 3), ms_fused_kernel_no_barriers=round(t_k, 3), ms_bcast_entry_local_destination_only=round(t_k1, 3), ms_local_gemm_256wide=round(t_k256, 3), tflops_fused=round(flop / t_f / 1e9, 1), tflops_nccl=round(flop / t_n / 1e9, 1),
          bytes_received_per_gpu=recv, nvlink_floor_ms_at_770GBs=round(recv / 770e9 * 1e3, 3), identical=bool(same))))
dist.destroy_process_group()
