"""Multi-GPU hot path: views are sharded for the cloud build, candidates are merged per track shard.

One process per GPU (torch.distributed, NCCL over NVLink).  Rank r unprojects a contiguous range of views and KEEPS
those clouds -- feature rows (512 B per point) never cross NVLink.  Per refinement iteration

    keys   = knn_keys(local clouds, all queries)            K1   [N, L, S, Ms, k] packed (d2, global index)
    recv   = all_to_all(keys)                                C1   rank j gets every rank's lists for track shard j
    idx_j  = topk_merge(recv)                                K2   exact global top-k for shard j
    idx    = all_gather(idx_j)                               C2
    out    = corr_from_idx(local clouds, idx)                K3   records of neighbours in local views, zeros elsewhere
    out    = all_reduce(out, SUM)                            C3   exactly one non-zero contributor per record

and every rank ends with the full [S, M, L*k*4] block (the transformer that consumes it is replicated and needs all
tracks: reference mvtracker/models/core/cotracker2/blocks.py:473-489).  Point indices are the reference's global ones
(v*H*W + y*W + x, model_utils.py:475), so results equal the single-GPU kernel's bit for bit in the neighbour sets and
offsets; correlations are the same fp32 sums.

north_star's literal plan (all-gather the clouds, then shard tracks) moves S*P*(C*e+16) bytes per window -- 8.6 GB
for the 8-view / 1M-point window -- which at NVLink speed costs as much as the whole single-GPU step; this design
moves ~1000x less ("Multi-GPU").  `backend` is the module providing the three kernels (mvtracker_b200.ops);
the CPU tests inject a numpy stand-in to exercise the exchanges over gloo.
"""
from typing import List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist


def _all_to_all(recv: torch.Tensor, send: torch.Tensor, group=None):
    """Equal-split all-to-all along dim 0.  NCCL: all_to_all_single.  gloo has no all-to-all (CPU tests only):
    all-gather and slice -- same result, plumbing only."""
    if dist.get_backend(group) != "gloo":
        dist.all_to_all_single(recv, send, group=group)
        return
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    everything = [torch.empty_like(send) for _ in range(world)]
    dist.all_gather(everything, send.contiguous(), group=group)
    for src in range(world):
        recv[src].copy_(everything[src][rank])


def view_range(rank: int, world: int, V: int) -> Tuple[int, int]:
    """Contiguous views [v0, v1) of `rank`; the first V % world ranks get one more.  Ranks beyond V get none."""
    base, rem = divmod(V, world)
    v0 = rank * base + min(rank, rem)
    return v0, v0 + base + (1 if rank < rem else 0)


def track_shard(world: int, M: int) -> int:
    """Tracks per shard (the last shard may be padded)."""
    return (M + world - 1) // world


class ShardedWindow:
    """One sliding window's clouds of THIS rank's views + the exchanges that make corr() global."""

    def __init__(self, clouds, v0: int, v1: int, V_total: int, k: int = 16, group=None, backend=None, n_levels=None):
        if backend is None:
            from . import ops as backend
        self.be, self.group = backend, group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.clouds, self.v0, self.v1, self.V_total, self.k = clouds, v0, v1, V_total, k
        self.levels = len(clouds) if n_levels is None else n_levels     # a rank without views still has to know
        self.width = self.levels * k * 4
        self._bufs = {}

    @classmethod
    def build(cls, fmaps_local, depths_local, intrs_local, extrs_local, v0: int, V_total: int, stride=4,
              levels: Sequence[int] = (0, 1, 2, 3), k=16, group=None, backend=None, reuse: Optional["ShardedWindow"] = None):
        """*_local: this rank's views only ([Vloc,S,C,H,W] ...); v0 = global index of the first one."""
        if backend is None:
            from . import ops as backend
        Vloc = fmaps_local.shape[0]
        clouds = []
        if Vloc > 0:
            clouds = backend.cloud_build(fmaps_local, depths_local, intrs_local, extrs_local, stride=stride, levels=levels,
                                         want_spacing=False, reuse=None if reuse is None else reuse.clouds)
        win = cls(clouds, v0, v0 + Vloc, V_total, k, group, backend, n_levels=len(list(levels)))
        if reuse is not None:
            win._bufs = reuse._bufs
        return win

    def _buf(self, name, shape, dtype, device):
        t = self._bufs.get(name)
        if t is None or tuple(t.shape) != tuple(shape) or t.dtype != dtype or t.device != device:
            t = torch.empty(shape, dtype=dtype, device=device)
            self._bufs[name] = t
        return t

    def corr(self, targets: torch.Tensor, coords: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """targets [S,M,C], coords [S,M,3] (replicated on every rank) -> [S, M, levels*k*4], identical on every rank."""
        be, N, k, L = self.be, self.world, self.k, self.levels
        S, M, _ = targets.shape
        Ms = track_shard(N, M)
        dev = targets.device
        keys = self._buf("keys", (N, L, S, Ms, k), torch.int64, dev)
        if self.clouds:
            offs = [self.v0 * c.H * c.W for c in self.clouds]
            be.knn_keys(self.clouds, offs, coords, k, n_shards=N, out=keys)
        else:
            keys.fill_(be.EMPTY_KEY)
        if N > 1:
            recv = self._buf("recv", (N, L, S, Ms, k), torch.int64, dev)
            _all_to_all(recv, keys, group=self.group)
        else:
            recv = keys
        idx_mine = self._buf("idx_mine", (L * S * Ms, k), torch.int32, dev)
        be.topk_merge(recv.view(N, L * S * Ms, k), out=idx_mine)
        idx_all = self._buf("idx_all", (N, L, S, Ms, k), torch.int32, dev)
        if N > 1:
            dist.all_gather_into_tensor(idx_all.view(N, -1), idx_mine.view(1, -1), group=self.group)
        else:
            idx_all.view(-1).copy_(idx_mine.view(-1))
        if out is None:
            out = torch.empty(S, M, L * k * 4, device=dev, dtype=torch.float32)
        if self.clouds:
            owned = [(c.xyz4, c.fvec, self.v0 * c.H * c.W, self.v1 * c.H * c.W) for c in self.clouds]
            be.corr_from_idx(owned, idx_all, targets, coords, k, out=out)
        else:
            out.zero_()
        if N > 1:
            dist.all_reduce(out, op=dist.ReduceOp.SUM, group=self.group)
        return out
