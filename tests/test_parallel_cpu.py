"""world_size-2 gloo tests (CPU) of the multi-GPU driver's host logic: shard bounds and the
assembly of M-sharded / N-sharded results.  The local contraction is injected (the oracle's fp32
matmul) because the product's local op is CUDA-only."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from tilelang_mesh_b200.parallel import ShardedGemm, all_shard_bounds, shard_bounds


def test_shard_bounds_cover_and_follow_ceil_rule():
    for dim in (0, 1, 7, 8, 65536, 1000):
        for world in (1, 2, 3, 4, 8):
            b = all_shard_bounds(dim, world)
            assert b[0][0] == 0 and b[-1][1] == dim
            assert all(b[i][1] == b[i + 1][0] for i in range(world - 1))
            per = -(-dim // world)
            assert all(hi - lo <= per for lo, hi in b)
    assert all_shard_bounds(65536, 8) == [(i * 8192, (i + 1) * 8192) for i in range(8)]  # C5 -> 8 x C2
    assert all_shard_bounds(10, 4) == [(0, 3), (3, 6), (6, 9), (9, 10)]
    with pytest.raises(ValueError):
        shard_bounds(8, 2, 2)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, mode, M, N, K, q):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import gemm_ref as R
        torch.manual_seed(0)
        A = torch.randint(-3, 4, (M, K)).float()
        B = torch.randint(-3, 4, (N, K)).float()
        want = A @ B.t()

        def local(a, b, c):
            c.copy_(R.matmul_fp32(a, b))

        sg = ShardedGemm(M, N, K, mode=mode, panel_rows=5, local_gemm=local)
        if mode == "M":
            a_loc = A[sg.lo:sg.hi]
            c_loc = sg.run(a_loc, B)
            ok = torch.equal(c_loc, want[sg.lo:sg.hi])
            full = sg.run(a_loc, B, gather=True)
            ok = ok and torch.equal(full, want)
        else:
            full = sg.run(A, B[sg.lo:sg.hi])
            ok = torch.equal(full, want)
        q.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("mode,M,N,K", [("M", 13, 6, 8), ("M", 16, 8, 8), ("N", 12, 8, 16), ("N", 7, 4, 8)])
def test_world2_gloo(mode, M, N, K):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, mode, M, N, K, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res == [(0, True), (1, True)]


def test_mode_n_needs_divisible_n(monkeypatch):
    monkeypatch.setattr(dist, "is_initialized", lambda: True)
    monkeypatch.setattr(dist, "get_world_size", lambda group=None: 2)
    monkeypatch.setattr(dist, "get_rank", lambda group=None: 1)
    with pytest.raises(ValueError, match="divisible"):
        ShardedGemm(8, 7, 8, mode="N", local_gemm=lambda a, b, c: None)
    sg = ShardedGemm(9, 8, 8, mode="M", local_gemm=lambda a, b, c: None)
    assert (sg.lo, sg.hi) == (5, 9) and sg.local_shapes() == dict(A=(4, 8), B=(8, 8), C=(4, 8))
