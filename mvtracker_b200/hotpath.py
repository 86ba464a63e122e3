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
                                 depth_interp_mode=depth_interp_mode, want_tiles=(search == "raster"))
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


class WindowSequence:
    """The sliding windows of ONE sequence with a per-frame cloud cache (exact): window w covers frames hop*w .. hop*w+S-1
    (mvtracker.py:522-566: S = 12, hop = S // 2) and shares S - hop frames with window w-1, whose clouds are kept in a
    CloudArena instead of being rebuilt (the reference rebuilds all S frames of every window, mvtracker.py:297-306).

        seq = WindowSequence(V, H', W', C, dtype, device)
        win = seq.push(fmaps_new, depths_new, intrs_win, extrs_win)      # first call: S frames; then `hop` new frames each
        fcorrs = win.corr(ffeats, coords)                                # a WindowClouds over the arena's current run

    Each push returns the WindowClouds of the new window; the one returned before stays valid until the run wraps around
    (slots / hop - 1 pushes later at the latest)."""

    def __init__(self, V: int, H: int, W: int, C: int, dtype, device, S: int = 12, hop: Optional[int] = None, stride=4,
                 levels: Sequence[int] = (0, 1, 2, 3), k=16, groups=1, add_offset=True, add_xyz=False, slots: Optional[int] = None,
                 want_valid: bool = False, depth_interp_mode="nearest"):
        self.S, self.hop = int(S), int(hop or S // 2)
        assert 0 < self.hop <= self.S
        if slots is None:       # up to four windows' worth of frames, within a third of the free memory
            per = ops.CloudArena.bytes_per_frame(V, H, W, C, torch.empty(0, dtype=dtype).element_size(), levels)
            free = torch.cuda.mem_get_info(device)[0]
            slots = int(max(self.S + self.hop, min(4 * self.S, (free // 3) // max(per, 1))))
        if slots < self.S + self.hop:
            raise ValueError(f"a sliding window of {self.S} frames advancing by {self.hop} needs at least {self.S + self.hop} slots")
        self.arena = ops.CloudArena(V, H, W, C, dtype, device, levels, slots, stride, want_valid=want_valid)
        self.args = (k, groups, add_offset, add_xyz)
        self.mode = depth_interp_mode
        self.slot0 = None           # first slot of the current window's run
        self.windows = 0
        self.frames_built = 0
        # wrap-around: the S - hop frames the NEXT window keeps are copied to the front of the arena on a side stream while the
        # current window's iterations run (they only read the run; the correlation kernel leaves most of the HBM bandwidth idle)
        self._side = torch.cuda.Stream(device=device) if torch.device(device).type == "cuda" else None
        self._moved = None          # event: the copy to the front has finished

    def push(self, fmaps_new, depths_new, intrs_win, extrs_win) -> WindowClouds:
        """fmaps_new [V, n, C, H', W'] / depths_new [V, n, H', W'] (or [V, n, 1, H', W']): the frames this window adds -- all S
        of the first window, `hop` afterwards (frame runs of larger tensors are taken as they are, no copy).
        intrs_win [V, S, 3, 3], extrs_win [V, S, 3, 4]: the cameras of the WHOLE window."""
        n, S, hop, ar = fmaps_new.shape[1], self.S, self.hop, self.arena
        if self.slot0 is None:
            if n != S:
                raise ValueError(f"the first window needs all {S} frames, got {n}")
            self.slot0, first_new = 0, 0
        else:
            if n != hop:
                raise ValueError(f"a later window adds {hop} frames, got {n}")
            self.slot0 += hop
            if self.slot0 + S > ar.slots:                      # wrap: keep the S - hop shared frames, at the front
                if self._moved is not None:                    # ... where the side stream has put them already
                    torch.cuda.current_stream().wait_event(self._moved)
                    self._moved = None
                else:
                    ar.move(self.slot0, 0, S - hop)
                self.slot0 = 0
            first_new = S - hop
        cam = ops.camera_prep(intrs_win, extrs_win)             # [V*S, CAM_STRIDE], record v*S + s
        cam_run = cam.reshape(-1)[first_new * cam.shape[1]:]
        ar.build(self.slot0 + first_new, fmaps_new, depths_new, cam_run, S, self.mode)
        self.windows += 1
        self.frames_built += n
        nxt = self.slot0 + hop
        if self._side is not None and nxt + S > ar.slots and self.slot0 >= S and not torch.cuda.is_current_stream_capturing():
            # the next push wraps: start moving the frames it keeps (this window's last S - hop) now; slot0 >= S: neither this
            # window's run nor the one returned before it touches the front slots
            self._side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(self._side):
                ar.move(nxt, 0, S - hop)
                self._moved = torch.cuda.Event()
                self._moved.record()
        return WindowClouds(ar.run(self.slot0, S, cam), None, *self.args)
