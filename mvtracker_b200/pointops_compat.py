"""pointops-shaped entry point: the one NATIVE precedent the reference has on this path is

    idx, dist = pointops.knn_query(k, xyz, offset, new_xyz, new_offset)        (mvtracker.py:55-65)

with `xyz [sum N_i, 3]` / `new_xyz [sum M_i, 3]` float32 contiguous, `offset` / `new_offset` the CUMULATIVE segment ends
(int), `idx [sum M_i, k]` int32 -- indices into the flattened `xyz`, i.e. GLOBAL, which is why _knn_pointops subtracts b*N
afterwards (mvtracker.py:68-69) -- and `dist [sum M_i, k]` float32 Euclidean, ascending.  pointops itself (third party,
ethz-vlg/pointcept@2082918 libs/pointops, not vendored, no network) cannot run here, so parity at this boundary is pinned
on the semantics the reference relies on: per-segment exact kNN in canonical (d2, index) order (tests/test_gpu_dropin.py).

Segments may have different sizes (pointops' general case; the reference happens to call it with equal ones): every
segment gets its own hashed grid (mvt_grid_build, up to 8 per launch set) and its own mvt_knn call.  The offsets are read on
the host (one small device->host copy): this is an interop entry, the hot path (dropin / hotpath) never goes through it.
"""
from typing import Tuple

import torch

from . import ops
from ._lib import MAX_JOBS, MvtError


def knn_query(k: int, xyz: torch.Tensor, offset: torch.Tensor, new_xyz: torch.Tensor = None,
              new_offset: torch.Tensor = None) -> Tuple[torch.Tensor, torch.Tensor]:
    """-> (idx [sum M, k] int32 global indices into xyz, -1 where a segment has fewer than k points; dist [sum M, k] fp32)."""
    if new_xyz is None:
        new_xyz, new_offset = xyz, offset
    if not xyz.is_cuda or not new_xyz.is_cuda:
        raise MvtError("mvtracker_b200.pointops_compat.knn_query needs CUDA tensors (no CPU fallback)")
    if xyz.dim() != 2 or xyz.shape[1] != 3 or new_xyz.dim() != 2 or new_xyz.shape[1] != 3:
        raise MvtError("xyz / new_xyz must be [n, 3]")
    ends, qends = [int(v) for v in offset.tolist()], [int(v) for v in new_offset.tolist()]
    if len(ends) != len(qends):
        raise MvtError(f"offset has {len(ends)} segments, new_offset {len(qends)}")
    if (ends and ends[-1] != xyz.shape[0]) or (qends and qends[-1] != new_xyz.shape[0]):
        raise MvtError("the last offset must equal the number of points")
    x, q = xyz.float().contiguous(), new_xyz.float().contiguous()
    idx = torch.full((q.shape[0], k), -1, device=q.device, dtype=torch.int32)
    dist = torch.full((q.shape[0], k), float("inf"), device=q.device, dtype=torch.float32)
    segs, a, qa = [], 0, 0
    for b, qb in zip(ends, qends):
        if b < a or qb < qa:
            raise MvtError("offsets must be non-decreasing")
        segs.append((a, b, qa, qb))
        a, qa = b, qb
    todo = [s for s in segs if s[1] > s[0] and s[3] > s[2]]
    for i in range(0, len(todo), MAX_JOBS):
        batch = todo[i:i + MAX_JOBS]
        grids = ops.build_grids([(x[a:b][None], None, None) for a, b, _, _ in batch], k_hint=k)
        for g, (a, b, qa, qb) in zip(grids, batch):
            d, ix = ops.knn(g, q[qa:qb][None], min(k, ops._lib.MAX_K), idx_dtype=torch.int32)
            kk = ix.shape[-1]
            idx[qa:qb, :kk] = torch.where(ix[0] >= 0, ix[0] + a, ix[0])
            dist[qa:qb, :kk] = d[0]
    return idx, dist
