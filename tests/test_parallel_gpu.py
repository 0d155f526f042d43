"""Multi-GPU driver on real GPUs (needs >= 2 devices; skipped otherwise): M-sharded (no collective)
and N-sharded (+ NCCL all-gather) contractions against the fp32 oracle of the full problem."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q, q_ipc):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        from oracle import gemm_ref as R
        from tilelang_mesh_b200 import gemm as G
        from tilelang_mesh_b200.parallel import ShardedGemm
        G._ensure_init(rank)
        # Predicate only, nothing is launched on these addresses: memory of another GPU, however it is mapped into
        # this process, must never be reported as this GPU's own (the condition for TMA multicast loads).
        shared = torch.zeros(1 << 20, dtype=torch.uint8, device="cuda")
        if rank == 1:
            q_ipc.put(shared)  # CUDA IPC: rank 0 maps GPU 1's allocation into its own address space
        else:
            imported = q_ipc.get(timeout=120)
            assert imported.device.index == 1
            assert not G.is_own_device_memory(imported.data_ptr(), imported.numel()), "IPC-imported peer memory"
            del imported
        dist.barrier()
        M, N, K = 4096, 2048, 1024
        torch.manual_seed(0)
        A = torch.randn(M, K, device="cuda").bfloat16()
        B = torch.randn(N, K, device="cuda").bfloat16()
        want = R.matmul_fp32(A, B).bfloat16()
        ok = True
        sg = ShardedGemm(M, N, K, mode="M")
        c_loc = sg.run(A[sg.lo:sg.hi].contiguous(), B)
        R.assert_allclose(c_loc, want[sg.lo:sg.hi], 1e-2, 1e-2)
        full = sg.run(A[sg.lo:sg.hi].contiguous(), B, gather=True)
        R.assert_allclose(full, want, 1e-2, 1e-2)
        sn = ShardedGemm(M, N, K, mode="N", panel_rows=1024)
        full_n = sn.run(A, B[sn.lo:sn.hi].contiguous())
        torch.cuda.synchronize()
        R.assert_allclose(full_n, want, 1e-2, 1e-2)
        # fused: the epilogue stores into every rank's C over NVLink (no NCCL on the data path)
        sf = ShardedGemm(M, N, K, mode="N", fused=True)
        for g_, ptr in enumerate(sf._symm_hdl.buffer_ptrs):  # symmetric memory: virtual-memory mappings, peers' and own
            assert not G.is_own_device_memory(int(ptr), 4096), f"symmetric-memory buffer of rank {g_}"
        for _ in range(2):
            full_f = sf.run(A, B[sf.lo:sf.hi].contiguous())
            torch.cuda.synchronize()
            ok = ok and torch.equal(full_f.view(torch.int16), full_n.view(torch.int16))
        # every rank holds the same gathered result
        ref = full_n.clone()
        dist.broadcast(ref, src=0)
        ok = ok and torch.equal(ref.view(torch.int16), full_n.view(torch.int16))
        q.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs at least 2 GPUs")
def test_sharded_gemm_two_gpus():
    ctx = mp.get_context("spawn")
    q, q_ipc = ctx.Queue(), ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q, q_ipc)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=300) for _ in procs)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert res == [(0, True), (1, True)]
