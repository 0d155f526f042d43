"""Multi-GPU driver for large contractions: one process per GPU over torch.distributed.

The reference has no multi-GPU runtime; its only sharding vocabulary is the mesh
annotation `T.MeshShardingPolicy` with `ceil(dim / G)` shard sizing
(tilelang/language/v2/annot.py:525-592).  The contraction shards by independent units:

 * mode "M" (primary): A and C split into G contiguous row blocks, B replicated.  Rows of C depend
   only on the matching rows of A, so there is NO data-path collective.
 * mode "N" (optional): B split into G row blocks (B is [N, K]), A replicated; every rank produces
   the column block C[:, n0:n1] and an all-gather over NCCL/NVLink materialises the full C on every
   rank.  The gather is issued per M-panel on a side stream so that the transfer of panel p overlaps
   the GEMM of panel p+1.  It is communication-bound by construction (each rank receives (G-1)/G of C).

The local contraction is the tcgen05 kernel through the C ABI (`gemm.gemm`); tests may inject another
`local_gemm` to exercise the sharding / gather logic on CPU with the gloo backend.
"""
from __future__ import annotations

from typing import Callable, List, Optional, Tuple

import torch
import torch.distributed as dist


def shard_bounds(dim: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous shard [lo, hi) of `dim` for `rank`: ceil(dim / world) per shard, last shards may be
    short or empty (annot.py:578-580 sizing)."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError(f"bad rank {rank} / world {world}")
    per = -(-dim // world)
    lo = min(rank * per, dim)
    return lo, min(lo + per, dim)


def all_shard_bounds(dim: int, world: int) -> List[Tuple[int, int]]:
    return [shard_bounds(dim, world, r) for r in range(world)]


class ShardedGemm:
    """C[M,N] = A[M,K] @ B[N,K]^T sharded over the ranks of `group`."""

    def __init__(self, M: int, N: int, K: int, mode: str = "M", group=None, panel_rows: int = 8192,
                 local_gemm: Optional[Callable] = None, fused: bool = False, dtype: torch.dtype = torch.bfloat16,
                 fused_cfg: Optional[dict] = None):
        if mode not in ("M", "N"):
            raise ValueError("mode must be 'M' (row-sharded A/C, no collective) or 'N' (row-sharded B + all-gather)")
        self.M, self.N, self.K, self.mode, self.group = M, N, K, mode, group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.panel_rows = max(1, panel_rows)
        if local_gemm is None:
            from . import gemm as G
            local_gemm = lambda a, b, c: G.gemm(a, b, c)  # noqa: E731
        self.local_gemm = local_gemm
        dim = M if mode == "M" else N
        self.lo, self.hi = shard_bounds(dim, self.world, self.rank)
        if mode == "N" and N % self.world:
            raise ValueError(f"mode 'N' needs N ({N}) divisible by the world size ({self.world}) for the all-gather")
        # fused = the GEMM kernel's epilogue stores every C tile into all ranks' copies of C over NVLink
        # (tlb200_gemm_bcast): compute and all-gather are one kernel, no NCCL call on the data path.
        self.fused = bool(fused) and mode == "N" and self.world > 1
        self.fused_cfg = dict(fused_cfg or {})  # plan hints of the fused kernel (e.g. epilogue=3: peer stores through st.global)
        self._symm_out = self._symm_hdl = None
        if self.fused:
            if self.world > 8:
                raise ValueError("fused all-gather supports at most 8 ranks (one NVSwitch domain)")
            import torch.distributed._symmetric_memory as symm_mem
            dev = torch.device("cuda", torch.cuda.current_device())
            self._symm_out = symm_mem.empty((M, N), dtype=dtype, device=dev)
            grp = group if group is not None else dist.group.WORLD
            self._symm_hdl = symm_mem.rendezvous(self._symm_out, grp)

    # shapes this rank must hold
    def local_shapes(self):
        r = self.hi - self.lo
        if self.mode == "M":
            return dict(A=(r, self.K), B=(self.N, self.K), C=(r, self.N))
        return dict(A=(self.M, self.K), B=(r, self.K), C=(self.M, r))

    def run(self, a: torch.Tensor, b: torch.Tensor, c: Optional[torch.Tensor] = None, gather: Optional[bool] = None):
        """Returns this rank's C block (mode "M", gather False) or the full C (gather True / mode "N")."""
        sh = self.local_shapes()
        if tuple(a.shape) != sh["A"] or tuple(b.shape) != sh["B"]:
            raise ValueError(f"rank {self.rank}: expected A{sh['A']} B{sh['B']}, got A{tuple(a.shape)} B{tuple(b.shape)}")
        gather = (self.mode == "N") if gather is None else gather
        if self.mode == "M":
            if c is None:
                c = torch.empty(sh["C"], dtype=a.dtype, device=a.device)
            if sh["C"][0] > 0:
                self.local_gemm(a, b, c)
            if not gather or self.world == 1:
                return c
            per = -(-self.M // self.world)
            full = torch.empty((per * self.world, self.N), dtype=c.dtype, device=c.device)
            pad = c if c.shape[0] == per else torch.cat([c, c.new_zeros(per - c.shape[0], self.N)])
            dist.all_gather_into_tensor(full, pad.contiguous(), group=self.group)
            return full[:self.M]
        return self._run_n_sharded(a, b)

    def _run_n_fused(self, a, b):
        from . import gemm as G
        out, hdl = self._symm_out, self._symm_hdl
        if a.dtype != out.dtype:
            raise ValueError(f"fused all-gather was set up for {out.dtype}, got {a.dtype}")
        off = self.lo * out.element_size()
        ptrs = [int(p) for p in hdl.buffer_ptrs]
        order = [self.rank] + [g for g in range(self.world) if g != self.rank]  # destination 0 = local copy
        hdl.barrier(channel=0)  # nobody may still be reading the previous result
        G.gemm_bcast(a, b, [ptrs[g] + off for g in order], ldc=self.N, c_dtype=out.dtype, **self.fused_cfg)
        hdl.barrier(channel=1)  # every rank's stores have landed everywhere
        return out

    def _run_n_sharded(self, a, b):
        if self.fused:
            return self._run_n_fused(a, b)
        G, ncol = self.world, self.hi - self.lo
        out = torch.empty((self.M, self.N), dtype=a.dtype, device=a.device)
        if G == 1:
            self.local_gemm(a, b, out)
            return out
        use_side = a.is_cuda
        comm = torch.cuda.Stream() if use_side else None
        pending = []
        for r0 in range(0, self.M, self.panel_rows):
            r1 = min(r0 + self.panel_rows, self.M)
            cpanel = torch.empty((r1 - r0, ncol), dtype=a.dtype, device=a.device)
            self.local_gemm(a[r0:r1], b, cpanel)
            stage = torch.empty((G, r1 - r0, ncol), dtype=a.dtype, device=a.device)
            if use_side:
                ev = torch.cuda.Event()
                ev.record()
                with torch.cuda.stream(comm):
                    comm.wait_event(ev)
                    dist.all_gather_into_tensor(stage.view(G * (r1 - r0), ncol), cpanel, group=self.group)
                    # [G, rows, N/G] -> out[rows, G*N/G]: column block g comes from rank g
                    out[r0:r1].view(r1 - r0, G, ncol).copy_(stage.permute(1, 0, 2))
                cpanel.record_stream(comm)
                stage.record_stream(comm)
            else:
                dist.all_gather_into_tensor(stage.view(G * (r1 - r0), ncol), cpanel, group=self.group)
                out[r0:r1].view(r1 - r0, G, ncol).copy_(stage.permute(1, 0, 2))
            pending.append((cpanel, stage))
        if use_side:
            torch.cuda.current_stream().wait_stream(comm)
        return out
