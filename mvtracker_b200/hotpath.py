"""The hot path of one sliding window as two calls (the loop structure of forward_iteration, reference
mvtracker/models/core/mvtracker/mvtracker.py:296-322 and :350-376, without the per-level Python round trips):

    win = WindowClouds.build(fmaps, depths, intrs, extrs)        # 4-level cloud in one pass; raster search: no index build
    fcorrs = win.corr(ffeats, coords)                            # all levels, all frames, all tracks: 1 launch
    # fcorrs [S, N, levels*k*(G+3)] == torch.cat([corr_sample(level l) ...], -1) at mvtracker.py:374
"""
from typing import List, Optional, Sequence

import torch

from . import ops


class WindowClouds:
    def __init__(self, clouds: List[ops.Cloud], grids: Optional[List[ops.Grid]], k: int, groups: int, add_offset: bool,
                 add_xyz: bool):
        self.clouds, self.grids = clouds, grids
        self.searches = ([ops.Search.from_cloud(c) for c in clouds] if grids is None
                         else [ops.Search.from_grid(g, c.xyz) for g, c in zip(grids, clouds)])
        self.k, self.groups, self.add_offset, self.add_xyz = k, groups, add_offset, add_xyz
        self.D = groups + 3 * bool(add_offset) + 3 * bool(add_xyz)
        self.width = len(clouds) * k * self.D
        # True: iteration i+1 starts its searches from iteration i's neighbours (their largest distance caps the k-th best;
        # exact either way).  Panoptic window, tracks moving N(0, 1 cm) per iteration: 0.697 -> 0.667 ms per warm iteration
        # (tools/diag_warm.py).  Off by default: bench.py times every iteration cold unless --warm-start is given.
        self.warm_start = False
        self._idx = None            # [levels, S, N, k] int32 of the last corr() call on these clouds

    @classmethod
    def build(cls, fmaps, depths, intrs, extrs, stride=4, levels: Sequence[int] = (0, 1, 2, 3), k=16, groups=1,
              add_offset=True, add_xyz=False, cell_scale=0.0, depth_interp_mode="nearest", search="raster") -> "WindowClouds":
        """fmaps [V,S,C,H',W'] (or with a leading batch axis of 1), depths [V,S,H',W'] / [V,S,1,H',W'].
        search: "raster" (search the depth rasters directly, nothing to build) or "grid" (hashed grid, 6 more launches)."""
        if fmaps.dim() == 6:
            assert fmaps.shape[0] == 1, "B must be 1 (mvtracker.py:275)"
            fmaps, depths, intrs, extrs = fmaps[0], depths[0], intrs[0], extrs[0]
        clouds = ops.cloud_build(fmaps, depths, intrs, extrs, stride=stride, levels=levels,
                                 depth_interp_mode=depth_interp_mode)
        grids = None
        if search == "grid":
            grids = ops.build_grids([(c.xyz, None, c.spacing) for c in clouds], k_hint=k, cell_scale=cell_scale)
        elif search != "raster":
            raise ValueError(search)
        return cls(clouds, grids, k, groups, add_offset, add_xyz)

    def corr(self, targets: torch.Tensor, coords: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """targets [S,N,C], coords [S,N,3] -> [S,N,width]."""
        levels = [(sr, c.fvec, self.k) for sr, c in zip(self.searches, self.clouds)]
        hint = idx = None
        if self.warm_start and self.grids is None:
            shape = (len(levels), targets.shape[0], targets.shape[1], self.k)
            if self._idx is not None and tuple(self._idx.shape) == shape:
                hint = idx = self._idx
            else:
                idx = self._idx = torch.empty(shape, device=targets.device, dtype=torch.int32)
        return ops.knn_corr(levels, targets, coords, groups=self.groups, add_offset=self.add_offset, add_xyz=self.add_xyz,
                            out=out, idx_hint=hint, idx_out=idx)

    def transformer_input(self, ffeats: torch.Tensor, coords: torch.Tensor, track_mask: torch.Tensor, vis: torch.Tensor,
                          pos_embed: Optional[torch.Tensor], time_embed: Optional[torch.Tensor], flow_embed_dim: int = 64,
                          out: Optional[torch.Tensor] = None) -> torch.Tensor:
        """One refinement iteration's transformer input x [N, S, D] (mvtracker.py:350-386) in two launches: the fused kNN
        correlation of all levels writes its block straight into x's correlation columns, mvt_track_input writes the
        flow embedding, the track features and the two masks and adds pos / time embeddings -- every byte of x is
        written once.  ffeats [S,N,C], coords [S,N,3], track_mask / vis [S,N]."""
        S, N, C = ffeats.shape
        D_n = self.groups + 3 * bool(self.add_offset) + 3 * bool(self.add_xyz)
        width = len(self.clouds) * self.k * D_n
        D = 3 * flow_embed_dim + 3 + width + C + 2
        if out is None:
            out = torch.empty(N, S, D, device=ffeats.device, dtype=torch.float32)
        self.corr(ffeats, coords, out=ops.corr_columns(out, flow_embed_dim, width))
        return ops.track_input(coords, ffeats, track_mask, vis, pos_embed, time_embed, flow_embed_dim, width, out=out,
                               corr_in_out=True)

    def feature_init(self, frame: int, query_xyz: torch.Tensor) -> torch.Tensor:
        """Nearest level-0 point's feature for each query of `frame` (mvtracker.py:627-643).  query_xyz [n,3] -> [n,C]."""
        c = self.clouds[0]
        if self.grids is None:
            sr = ops.Search.from_cloud(c, frame=frame, tiles=True)
            return ops.knn_gather(sr, c.fvec[frame:frame + 1], query_xyz[None])[0]
        g = ops.build_grids([(c.xyz[frame:frame + 1], None, c.spacing[frame:frame + 1])], k_hint=1)[0]
        return ops.knn_gather(g, c.fvec[frame:frame + 1], query_xyz[None])[0]
