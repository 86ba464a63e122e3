"""Torch-facing wrappers over the C ABI (include/mvt.h).  PyTorch is plumbing only here: it owns device memory and the
current stream; every computation below runs in libmvt.so's sm_100a kernels.  CUDA tensors only -- there is no CPU path."""
import os
from dataclasses import dataclass
from typing import List, Optional, Sequence

import torch

from . import _lib
from ._lib import C, CorrLevel, GridJob, LevelOut, MvtError, check, lib
from ._lib import Search as _CSearch


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _need_cuda(t: torch.Tensor, name: str):
    if not isinstance(t, torch.Tensor) or not t.is_cuda:
        raise MvtError(f"{name} must be a CUDA tensor: mvtracker_b200 has no CPU fallback (use the reference on CPU)")


def _ptr(t: Optional[torch.Tensor]):
    return C.c_void_p(None) if t is None else C.c_void_p(t.data_ptr())


def _feat_dtype(t: torch.Tensor) -> int:
    if t.dtype == torch.float32:
        return _lib.MVT_F32
    if t.dtype == torch.bfloat16:
        return _lib.MVT_BF16
    raise MvtError(f"feature dtype {t.dtype} not supported (float32 or bfloat16)")


def camera_prep(intrs: torch.Tensor, extrs: torch.Tensor) -> torch.Tensor:
    """intrs [...,3,3], extrs [...,3,4] -> cam [n, CAM_STRIDE] fp32 records, see mvt_camera_prep."""
    _need_cuda(intrs, "intrs"); _need_cuda(extrs, "extrs")
    K = intrs.reshape(-1, 3, 3).float().contiguous()
    E = extrs.reshape(-1, 3, 4).float().contiguous()
    assert K.shape[0] == E.shape[0]
    cam = torch.empty(K.shape[0], _lib.CAM_STRIDE, device=K.device, dtype=torch.float32)
    check(lib().mvt_camera_prep(_ptr(K), _ptr(E), K.shape[0], _ptr(cam), _stream()), "mvt_camera_prep")
    return cam


@dataclass
class Cloud:
    """One pyramid level of the fused cloud: the reference's (pointcloud_xyz, pointcloud_fvec[, valid]) triple."""
    level: int
    xyz: torch.Tensor                  # [S, P, 3] fp32
    fvec: Optional[torch.Tensor]       # [S, P, C]
    valid: Optional[torch.Tensor]      # [S, P] bool
    spacing: Optional[torch.Tensor]    # [S, 2] (sum, count) footprint hint for the grid
    xyz4: Optional[torch.Tensor] = None  # [S, P, 4] {x,y,z,depth} raster records (the raster search structure)
    cam: Optional[torch.Tensor] = None   # [V*S, CAM_STRIDE] camera records
    V: int = 0
    H: int = 0                           # raster height / width of this level
    W: int = 0
    stride: float = 0.0                  # input pixels per raster pixel at this level
    tile_z: Optional[torch.Tensor] = None  # [S, V, ceil(H/8), ceil(W/8), 2] per-tile depth range (tile_depth_build)
    tiles_ready: bool = False              # tile_z describes the current xyz4 (a reused buffer starts out stale)
    neg: Optional[torch.Tensor] = None     # [S] int32: frame holds points with depth < 0 (mvt_level_out.neg_depth)


def tile_depth_build(clouds: Sequence["Cloud"], cam_view_stride: int = 0) -> None:
    """Fills cloud.tile_z of every Cloud that has raster records (one launch for all): the per-tile depth ranges that let
    the raster search skip, inside a large pixel box, the tiles the query's ball cannot reach (mvt_tile_depth_build)."""
    todo = [c for c in clouds if c.xyz4 is not None and c.cam is not None]
    if not todo or os.environ.get("MVT_NO_TILES"):       # (debugging knob: searches then read every pixel of a box)
        return
    if len(todo) > 8:
        raise MvtError("tile_depth_build takes at most 8 clouds")
    S = todo[0].xyz4.shape[0]
    arr = (_CSearch * len(todo))()
    outs = (C.c_void_p * len(todo))()
    for i, c in enumerate(todo):
        if c.xyz4.shape[0] != S:
            raise MvtError("tile_depth_build: clouds of one call must have the same number of frames")
        shape = (S, c.V, (c.H + 7) // 8, (c.W + 7) // 8, 2)
        if c.tile_z is None or tuple(c.tile_z.shape) != shape:
            c.tile_z = torch.empty(shape, device=c.xyz4.device, dtype=torch.float32)
        arr[i] = _CSearch(None, None, c.xyz4.data_ptr(), c.cam.data_ptr(), int(cam_view_stride) or S, c.xyz4.shape[1], c.V, c.H, c.W,
                          float(c.stride), 0, None, None)
        outs[i] = c.tile_z.data_ptr()
    check(lib().mvt_tile_depth_build(arr, len(todo), S, outs, _stream()), "mvt_tile_depth_build")
    for c in todo:
        c.tiles_ready = True


def _view_strided(t: torch.Tensor, inner: int):
    """t [V, S, ...] whose views are whole contiguous runs of S frames (inner elements per frame), any distance apart -- e.g.
    frames 6..11 of a [V,12,C,H,W] tensor.  -> (tensor to pass, elements between views); copies only when the layout is not that."""
    V, S = t.shape[0], t.shape[1]
    want = [inner]
    for d in range(t.dim() - 1, 1, -1):
        want.append(want[-1] // t.shape[d] if t.shape[d] else 0)
    ok = t.stride(1) == inner and (V == 1 or t.stride(0) >= S * inner)
    st, run = 1, True
    for d in range(t.dim() - 1, 1, -1):
        run = run and (t.shape[d] == 1 or t.stride(d) == st)
        st *= t.shape[d]
    if ok and run:
        return t, (t.stride(0) if V > 1 else S * inner)
    t = t.contiguous()
    return t, S * inner


def cloud_build(fmaps, depths, intrs, extrs, stride=4, levels: Sequence[int] = (0,), depth_interp_mode="nearest",
                want_valid=False, want_fvec=True, want_xyz4=True, want_spacing=True, cam=None,
                reuse: Optional[Sequence["Cloud"]] = None, shared_records=None, view_offset: int = 0,
                V_total: Optional[int] = None, want_tiles=False, cam_view_stride: int = 0) -> List[Cloud]:
    """fmaps [V,S,C,H,W], depths [V,S,H,W] (or [V,S,1,H,W]) -> one Cloud per requested level, built in ONE kernel pass.
    fmaps / depths may be frame runs of larger tensors (views any distance apart, e.g. x[:, 6:]): no copy is made.
    reuse: Clouds of an earlier call with the same shapes / levels / dtype whose buffers are overwritten in place (no
    allocation in steady state, e.g. one set of buffers per sliding window; CloudArena hands out frame runs of its slots).
    cam + cam_view_stride: camera records of a longer window, cam pointing at this run's first frame (records of
    consecutive views cam_view_stride apart); default: the records are made here from intrs / extrs.
    shared_records {level: _lib.Peers}: also store the 16-byte records of these views into the replicated
    [S, V_total*H_l*W_l, 4] buffers of every rank, views view_offset.. (mvt_cloud_build_shared).
    want_tiles: also build the per-tile depth tables now, in one launch (otherwise the first search-only call on a level
    builds that level's table; the fused knn_corr kernel uses them to hand its large-box queries to the tile search)."""
    _need_cuda(fmaps, "fmaps"); _need_cuda(depths, "depths")
    V, S, Cc, H, W = fmaps.shape
    depths = depths.reshape(V, S, H, W) if depths.dim() == 4 or depths.is_contiguous() else depths[:, :, 0]
    if depth_interp_mode not in ("nearest", "avg"):
        raise NotImplementedError(depth_interp_mode)
    levels = sorted(set(int(l) for l in levels))
    if not levels or levels[0] < 0 or levels[-1] >= _lib.MAX_LEVELS:
        raise MvtError(f"levels {levels} outside 0..{_lib.MAX_LEVELS - 1}")
    fm, fm_vs = _view_strided(fmaps, Cc * H * W)
    dp, dp_vs = _view_strided(depths if depths.dtype == torch.float32 else depths.float(), H * W)
    if reuse is not None:
        reuse = {c.level: c for c in reuse}
        if cam is None:
            cam = next(iter(reuse.values())).cam
            K = intrs.reshape(-1, 3, 3).float().contiguous()
            E = extrs.reshape(-1, 3, 4).float().contiguous()
            if cam is None or cam.shape[0] != K.shape[0]:
                raise MvtError("reuse: camera record buffer has the wrong size")
            check(lib().mvt_camera_prep(_ptr(K), _ptr(E), K.shape[0], _ptr(cam), _stream()), "mvt_camera_prep")
    if cam is None:
        cam = camera_prep(intrs, extrs)
    dev = fm.device
    outs = (LevelOut * _lib.MAX_LEVELS)()
    clouds, mask = [], 0
    for l in levels:
        Hl, Wl = H >> l, W >> l
        if Hl <= 0 or Wl <= 0:
            raise MvtError(f"level {l} is empty for a {H}x{W} map")
        P = V * Hl * Wl
        if reuse is not None:
            old = reuse.get(l)
            if old is None or old.xyz.shape != (S, P, 3) or (want_fvec and (old.fvec is None or old.fvec.shape != (S, P, Cc)
                                                                          or old.fvec.dtype != fm.dtype)):
                raise MvtError(f"reuse: no matching buffers for level {l}")
            xyz, fvec = old.xyz, (old.fvec if want_fvec else None)
            valid = old.valid.view(torch.uint8) if (want_valid and old.valid is not None) else None
            spacing = old.spacing if want_spacing else None
            xyz4 = old.xyz4 if want_xyz4 else None
            neg = old.neg if (want_xyz4 and old.neg is not None) else (torch.empty(S, device=dev, dtype=torch.int32) if want_xyz4 else None)
            if (want_valid and valid is None) or (want_spacing and spacing is None) or (want_xyz4 and xyz4 is None):
                raise MvtError(f"reuse: level {l} lacks a requested output buffer")
        else:
            xyz = torch.empty(S, P, 3, device=dev, dtype=torch.float32)
            fvec = torch.empty(S, P, Cc, device=dev, dtype=fm.dtype) if want_fvec else None
            valid = torch.empty(S, P, device=dev, dtype=torch.uint8) if want_valid else None
            spacing = torch.empty(S, 2, device=dev, dtype=torch.float32) if want_spacing else None
            xyz4 = torch.empty(S, P, 4, device=dev, dtype=torch.float32) if want_xyz4 else None
            neg = torch.empty(S, device=dev, dtype=torch.int32) if want_xyz4 else None
        outs[l] = LevelOut(xyz.data_ptr(), xyz4.data_ptr() if want_xyz4 else None, fvec.data_ptr() if want_fvec else None,
                           valid.data_ptr() if want_valid else None, spacing.data_ptr() if want_spacing else None,
                           neg.data_ptr() if neg is not None else None)
        mask |= 1 << l
        clouds.append(Cloud(l, xyz, fvec, valid.view(torch.bool) if want_valid else None, spacing, xyz4, cam, V, Hl, Wl,
                            float(stride) * (1 << l), old.tile_z if reuse is not None else None, False, neg))
    dense = fm_vs == S * Cc * H * W and dp_vs == S * H * W and not cam_view_stride
    recs = None
    if shared_records is not None:
        recs = (_lib.Peers * _lib.MAX_LEVELS)()
        for l in levels:
            recs[l] = shared_records[l]
    if recs is None and dense:
        check(lib().mvt_cloud_build(_ptr(fm), _feat_dtype(fm), _ptr(dp), _ptr(cam), V, S, Cc, H, W, float(stride),
                                    int(depth_interp_mode == "avg"), outs, mask, _stream()), "mvt_cloud_build")
    elif dense:
        check(lib().mvt_cloud_build_shared(_ptr(fm), _feat_dtype(fm), _ptr(dp), _ptr(cam), V, S, Cc, H, W, float(stride),
                                           int(depth_interp_mode == "avg"), outs, mask, recs, int(view_offset),
                                           int(V if V_total is None else V_total), _stream()), "mvt_cloud_build_shared")
    else:
        check(lib().mvt_cloud_build_frames(_ptr(fm), int(fm_vs), _feat_dtype(fm), _ptr(dp), int(dp_vs), _ptr(cam),
                                           int(cam_view_stride), V, S, Cc, H, W, float(stride), int(depth_interp_mode == "avg"),
                                           outs, mask, recs, int(view_offset), int(V if V_total is None else V_total), _stream()),
              "mvt_cloud_build_frames")
    if want_tiles and want_xyz4:
        tile_depth_build(clouds, cam_view_stride=cam_view_stride)
    return clouds


class CloudArena:
    """Per-level cloud arrays with room for `slots` frames of ONE sequence: the per-frame cloud cache of sliding windows.

    The reference rebuilds all S = 12 frames of every window although consecutive windows share S/2 of them
    (mvtracker.py:297-306 vs :549-555).  Here a window is a contiguous run of S slots; advancing by `hop` frames builds only
    the new frames into the slots right behind the run (mvt_cloud_build_frames), so every window's clouds are plain
    contiguous views [S, P, .] -- the kernels need no indirection.  When the run reaches the end of the arena the S - hop
    frames it keeps are copied to the front (once every (slots - S) / hop + 1 windows).  HBM is what pays for it: a
    panoptic frame (1.36 M points, fp32 features) is 0.74 GB, the default 4 windows' worth of slots 35 GB of the 180."""

    def __init__(self, V: int, H: int, W: int, C: int, dtype, device, levels: Sequence[int] = (0, 1, 2, 3), slots: int = 48,
                 stride: float = 4.0, want_valid: bool = True, want_xyz: bool = True):
        self.V, self.H, self.W, self.C, self.dtype, self.device = V, H, W, C, dtype, device
        self.levels, self.slots, self.stride = [int(l) for l in levels], int(slots), float(stride)
        self.buf = {}
        for l in self.levels:
            Hl, Wl = H >> l, W >> l
            P = V * Hl * Wl
            self.buf[l] = dict(
                xyz=torch.empty(slots, P, 3, device=device, dtype=torch.float32) if want_xyz else None,
                xyz4=torch.empty(slots, P, 4, device=device, dtype=torch.float32),
                fvec=torch.empty(slots, P, C, device=device, dtype=dtype),
                valid=torch.empty(slots, P, device=device, dtype=torch.uint8) if want_valid else None,
                neg=torch.zeros(slots, device=device, dtype=torch.int32),
                tile_z=torch.empty(slots, V, (Hl + 7) // 8, (Wl + 7) // 8, 2, device=device, dtype=torch.float32))

    @staticmethod
    def bytes_per_frame(V, H, W, C, esize, levels=(0, 1, 2, 3)) -> int:
        return sum(V * (H >> l) * (W >> l) * (12 + 16 + C * esize + 1) + 8 * V * (((H >> l) + 7) // 8) * (((W >> l) + 7) // 8)
                   for l in levels)

    def run(self, slot0: int, n: int, cam: Optional[torch.Tensor] = None, tiles_ready: bool = True) -> List[Cloud]:
        """Clouds over slots slot0 .. slot0+n-1 (views of the arena; cam: the camera records [V*n, CAM_STRIDE] of that run)."""
        assert 0 <= slot0 and slot0 + n <= self.slots
        out = []
        for l in self.levels:
            b = self.buf[l]
            sl = slice(slot0, slot0 + n)
            out.append(Cloud(l, b["xyz"][sl] if b["xyz"] is not None else None, b["fvec"][sl],
                             b["valid"][sl].view(torch.bool) if b["valid"] is not None else None, None, b["xyz4"][sl], cam,
                             self.V, self.H >> l, self.W >> l, self.stride * (1 << l), b["tile_z"][sl], tiles_ready, b["neg"][sl]))
        return out

    def build(self, slot0: int, fmaps: torch.Tensor, depths: torch.Tensor, cam_run: torch.Tensor, cam_view_stride: int,
              depth_interp_mode="nearest") -> None:
        """Builds the frames of fmaps [V, n, C, H, W] (a frame run of any tensor) into slots slot0 .. slot0+n-1, all levels and
        their tile tables.  cam_run: the window's camera records starting at this run's first frame (a view into the
        window's [V*S, CAM_STRIDE] record array), consecutive views cam_view_stride records apart."""
        n = fmaps.shape[1]
        tgt = self.run(slot0, n, cam_run, tiles_ready=False)
        for c in tgt:
            if c.xyz is None:           # cloud_build always writes xyz; an arena without it hands out the xyz4 records only
                raise MvtError("CloudArena(want_xyz=False) cannot be built into: the kernels write xyz")
        cloud_build(fmaps, depths, None, None, stride=self.stride, levels=self.levels, depth_interp_mode=depth_interp_mode,
                    want_valid=self.buf[self.levels[0]]["valid"] is not None, want_spacing=False, cam=cam_run, reuse=tgt,
                    want_tiles=True, cam_view_stride=cam_view_stride)

    def move(self, src: int, dst: int, n: int) -> None:
        """Copies n frames' clouds from slot src to slot dst (non-overlapping; the wrap-around of a sliding window)."""
        assert src >= dst + n or dst >= src + n
        for b in self.buf.values():
            for t in b.values():
                if t is not None:
                    t[dst:dst + n].copy_(t[src:src + n])


class Grid:
    """Search structure over B clouds of N points (built by build_grids)."""

    def __init__(self, xyz: torch.Tensor, mask: Optional[torch.Tensor], index_mode: int):
        self.xyz = xyz
        self.mask = mask
        self.index_mode = index_mode
        self.B, self.N = xyz.shape[0], xyz.shape[1]
        nbytes = lib().mvt_grid_workspace_bytes(self.B, self.N)
        self.workspace = torch.empty(nbytes + 256, device=xyz.device, dtype=torch.uint8)
        off = (-self.workspace.data_ptr()) % 256
        self.ptr = self.workspace.data_ptr() + off

    def info(self) -> torch.Tensor:
        out = torch.empty(self.B, 16, device=self.xyz.device, dtype=torch.float32)
        check(lib().mvt_grid_info(C.c_void_p(self.ptr), self.B, self.N, _ptr(out), _stream()), "mvt_grid_info")
        return out


def build_grids(clouds: Sequence, k_hint=16, index_mode=0, cell_scale=0.0) -> List[Grid]:
    """clouds: sequence of (xyz [B,N,3], mask [B,N] | None, spacing [B,2] | None).  One set of launches for all (<= 8)."""
    grids, jobs = [], (GridJob * _lib.MAX_JOBS)()
    if not 1 <= len(clouds) <= _lib.MAX_JOBS:
        raise MvtError(f"build_grids takes 1..{_lib.MAX_JOBS} clouds")
    keep = []
    for i, (xyz, mask, spacing) in enumerate(clouds):
        _need_cuda(xyz, "xyz")
        if xyz.dim() != 3 or xyz.shape[-1] != 3:
            raise MvtError(f"xyz must be [B,N,3], got {tuple(xyz.shape)}")
        x = xyz if (xyz.dtype == torch.float32 and xyz.is_contiguous()) else xyz.float().contiguous()
        m = None
        if mask is not None:
            m = mask.contiguous().view(torch.uint8) if mask.dtype == torch.bool else mask.to(torch.uint8).contiguous()
        g = Grid(x, m, index_mode)
        keep.append((x, m, spacing))
        jobs[i] = GridJob(x.data_ptr(), m.data_ptr() if m is not None else None,
                          spacing.data_ptr() if spacing is not None else None, g.ptr, g.B, g.N, int(k_hint), int(index_mode),
                          float(cell_scale))
        grids.append(g)
    check(lib().mvt_grid_build(jobs, len(clouds), _stream()), "mvt_grid_build")
    return grids


class Search:
    """What a query kernel searches in (mvt_search): a hashed grid over any cloud, or the raster records of a Cloud."""

    def __init__(self, c_struct, B, N, keep):
        self.c, self.B, self.N, self._keep = c_struct, B, N, keep

    @classmethod
    def from_grid(cls, grid: Grid, xyz: Optional[torch.Tensor] = None):
        x = grid.xyz if xyz is None else xyz
        c = _CSearch(grid.ptr, x.data_ptr(), None, None, 0, grid.N, 0, 0, 0, 0.0, 0, None, None)
        return cls(c, grid.B, grid.N, (grid, x))

    @classmethod
    def from_cloud(cls, cloud: Cloud, skip_invalid=False, frame: Optional[int] = None, nframes: int = 1, tiles=False):
        """Raster search over a Cloud from cloud_build (no build step).  frame=f restricts to frames f..f+nframes-1.
        tiles: the consumer is a search-only kernel (knn, knn_gather, knn_keys, knn_idx_shared), which culls large pixel
        boxes by the per-tile depth tables: build them now if this cloud's are missing or stale."""
        if tiles and cloud.xyz4 is not None and cloud.cam is not None and not cloud.tiles_ready:
            tile_depth_build([cloud])
        if cloud.xyz4 is None or cloud.cam is None:
            raise MvtError("this Cloud has no raster records (build it with want_xyz4=True)")
        S, P = cloud.xyz4.shape[0], cloud.xyz4.shape[1]
        x4, cam, B = cloud.xyz4, cloud.cam, S
        x4_ptr, cam_ptr = x4.data_ptr(), cam.data_ptr()
        if frame is not None:
            x4_ptr += frame * P * 16
            cam_ptr += frame * _lib.CAM_STRIDE * 4
            B = int(nframes)
            assert 0 <= frame and frame + B <= S
        tz = cloud.tile_z if cloud.tiles_ready else None
        tz_ptr = None
        if tz is not None:
            tz_ptr = tz.data_ptr() + (0 if frame is None else frame * tz[0].numel() * 4)
        neg_ptr = None
        if cloud.neg is not None:
            neg_ptr = cloud.neg.data_ptr() + (0 if frame is None else frame * 4)
        c = _CSearch(None, None, x4_ptr, cam_ptr, S, P, cloud.V, cloud.H, cloud.W, float(cloud.stride), int(bool(skip_invalid)),
                     tz_ptr, neg_ptr)
        return cls(c, B, P, (x4, cam, tz, cloud.neg))


def _as_search(obj, tiles=False) -> Search:
    if isinstance(obj, Search):
        return obj
    if isinstance(obj, Grid):
        return Search.from_grid(obj)
    if isinstance(obj, Cloud):
        return Search.from_cloud(obj, tiles=tiles)
    raise MvtError(f"cannot search in {type(obj)}")


def _f32c(t: torch.Tensor) -> torch.Tensor:
    return t if (t.dtype == torch.float32 and t.is_contiguous()) else t.float().contiguous()


def knn(where, query: torch.Tensor, k: int, idx_dtype=torch.int64, want_dist=True):
    """where: Grid | Cloud | Search.  -> (dists [B,M,k] fp32 ascending, idx [B,M,k]); canonical (d2, index) order;
    -1 / inf where fewer than k searchable points exist."""
    _need_cuda(query, "query")
    sr = _as_search(where, tiles=True)
    q = _f32c(query)
    B, M, _ = q.shape
    if B != sr.B:
        raise MvtError(f"query batch {B} != search batch {sr.B}")
    idx = torch.empty(B, M, k, device=q.device, dtype=idx_dtype)
    dist = torch.empty(B, M, k, device=q.device, dtype=torch.float32) if want_dist else None
    check(lib().mvt_knn(C.byref(sr.c), B, _ptr(q), M, int(k), _ptr(idx), int(idx_dtype == torch.int64), _ptr(dist),
                        _stream()), "mvt_knn")
    return dist, idx


_corr_ws = {}       # device index -> uint8 scratch of mvt_knn_corr (deferred-query list); grown on demand, reused by every call


def _corr_workspace(dev: torch.device, n_levels: int, B: int, M: int) -> torch.Tensor:
    need = int(lib().mvt_knn_corr_workspace_bytes(n_levels, B, M))
    ws = _corr_ws.get(dev.index)
    if ws is None or ws.numel() < need:
        ws = torch.empty(need + (need >> 2), device=dev, dtype=torch.uint8)      # calls of one device run on one stream at a time
        _corr_ws[dev.index] = ws
    return ws


class _HardQueryPolicy:
    """defer="auto": whether the next fused launch hands its large-box queries to the tile-search kernel.  Deferral pays when
    several per cent of the queries float in free space (a window that starts with static tracks, mvtracker.py:510) and costs
    ~8 % when only a few do; even COUNTING them costs ~3 % (the deferral build of the kernel is larger, and the fused kernel is
    sensitive to its instruction footprint).  So while few queries are hard the plain kernel runs (mode 0) and every PROBE-th
    launch counts (mode 2); while deferring (mode 1) every launch leaves its count anyway.  A count is copied to pinned host
    memory asynchronously and looked at when a later call is issued (no synchronisation: the decision is a call or more stale)."""
    ON, OFF = 0.015, 0.0075      # fraction of a launch's queries; hysteresis
    PROBE = 8

    def __init__(self):
        self.deferring = True    # a sequence starts with static tracks
        self.pinned = torch.zeros(1, dtype=torch.int32).pin_memory()
        self.event, self.queries, self.last_fraction = None, 0, None
        self.calls = 0

    def mode(self) -> int:
        if torch.cuda.is_current_stream_capturing():     # a captured launch keeps the mode it was captured with
            return 1 if self.deferring else 0
        if self.event is not None and self.event.query():
            self.last_fraction = float(self.pinned[0]) / max(1, self.queries)
            self.event = None
            if self.last_fraction > self.ON:
                self.deferring = True
            elif self.last_fraction < self.OFF:
                self.deferring = False
        if self.deferring:
            return 1
        return 2 if self.calls % self.PROBE == 0 else 0

    def record(self, ws: torch.Tensor, queries: int, mode: int):
        self.calls += 1
        # (mode 1: a sample every few launches -- the 4-byte copy costs a few microseconds of stream time)
        if mode != 0 and self.event is None and (mode == 2 or (self.calls & 3) == 1) and not torch.cuda.is_current_stream_capturing():
            self.pinned.copy_(ws[:4].view(torch.int32), non_blocking=True)
            self.queries = queries
            self.event = torch.cuda.Event()
            self.event.record()


_hard_policy = {}     # device index -> _HardQueryPolicy


def hard_query_policy(dev: torch.device) -> _HardQueryPolicy:
    pol = _hard_policy.get(dev.index)
    if pol is None:
        pol = _hard_policy[dev.index] = _HardQueryPolicy()
    return pol


def knn_corr(levels: Sequence, targets: torch.Tensor, coords: torch.Tensor, groups=1, add_offset=True, add_xyz=False,
             out: Optional[torch.Tensor] = None, return_idx=False, idx_hint: Optional[torch.Tensor] = None,
             idx_out: Optional[torch.Tensor] = None, defer="auto"):
    """levels: sequence of (where, fvec [B,N,C], k) with where = Grid | Cloud | Search.
    -> out [B, M, sum_l k_l*D] fp32 (level-major columns, the layout of torch.cat(fcorrs, -1) at mvtracker.py:374),
    D = groups + 3*add_offset + 3*add_xyz.
    idx_hint [L,B,M,k] int32: warm start (any k points per query, e.g. the previous iteration's idx); idx_out: buffer that
    receives the neighbour indices (may be the same tensor as idx_hint).
    defer: True = queries with large pixel boxes (tracks in free space) are finished by the tile-search kernel instead of
    holding the fused kernel's warps (needs tile tables on every level: cloud_build(..., want_tiles=True) or
    tile_depth_build); False = one kernel; "auto" = decided from the count the previous launch left (_HardQueryPolicy).
    Same results either way."""
    _need_cuda(targets, "targets"); _need_cuda(coords, "coords")
    B, M, Cc = targets.shape
    tg = targets.contiguous()
    cq = _f32c(coords)
    D = groups + 3 * bool(add_offset) + 3 * bool(add_xyz)
    n = len(levels)
    lv = (CorrLevel * _lib.MAX_LEVELS)()
    off, keep = 0, []
    for i, (where, fvec, k) in enumerate(levels):
        sr = _as_search(where)
        keep.append(sr)
        if sr.B != B:
            raise MvtError(f"targets batch {B} != search batch {sr.B}")
        if fvec.dtype != tg.dtype:
            raise MvtError(f"targets dtype {tg.dtype} != fvec dtype {fvec.dtype}")
        if not fvec.is_contiguous() or fvec.shape[-1] != Cc:
            raise MvtError("fvec must be contiguous [B,N,C] with the targets' C")
        lv[i] = CorrLevel(sr.c, fvec.data_ptr(), int(k), off)
        off += int(k) * D
    if out is None:
        out = torch.empty(B, M, off, device=tg.device, dtype=torch.float32)
    # out may be any [B, M, width] view with unit stride along the last axis, e.g. the column slice
    # tin.permute(1, 0, 2)[:, :, c0:c0 + width] of a [M, B, D_in] transformer-input buffer
    row_stride, frame_stride = out.stride(1), out.stride(0)
    assert out.shape[0] == B and out.shape[1] == M and out.shape[2] >= off and (out.stride(2) == 1 or out.shape[2] == 1)
    assert out.dtype == torch.float32
    idx = idx_out
    if return_idx and idx is None:
        idx = torch.empty(n, B, M, int(levels[0][2]), device=tg.device, dtype=torch.int32)
    for t in (idx, idx_hint):
        if t is not None:
            assert t.dtype == torch.int32 and t.is_contiguous() and t.numel() == n * B * M * int(levels[0][2])
    ws, mode, pol = None, 0, None
    if defer and not os.environ.get("MVT_NO_DEFER") and all(sr.c.tile_z for sr in keep):
        ws = _corr_workspace(tg.device, n, B, M)
        mode = 1
        if defer == "auto":
            pol = hard_query_policy(tg.device)
            mode = pol.mode()
    check(lib().mvt_knn_corr(lv, n, B, M, Cc, int(groups), _feat_dtype(tg), _ptr(tg), _ptr(cq), int(bool(add_offset)),
                             int(bool(add_xyz)), _ptr(out), row_stride, frame_stride, _ptr(idx_hint), _ptr(idx), _ptr(ws), mode,
                             _stream()), "mvt_knn_corr")
    if pol is not None:
        pol.record(ws, n * B * M, mode)
    return (out, idx) if return_idx else out


def offset_peers(peers: "_lib.Peers", nbytes: int) -> "_lib.Peers":
    """The same symmetric buffer, every mapping (and the multicast address) advanced by nbytes."""
    q = _lib.Peers()
    q.n = peers.n
    q.mc = (peers.mc + nbytes) if peers.mc else None
    for r in range(peers.n):
        q.ptr[r] = peers.ptr[r] + nbytes
    return q


class CorrPlan:
    """The per-level descriptors of a fused correlation call, made once for a set of clouds and reused by every launch on
    them (a refinement loop calls the kernel `iters` times per window on the same clouds: the ctypes structs are not rebuilt)."""

    def __init__(self, levels: Sequence):
        self.n = len(levels)
        self.lv = (CorrLevel * _lib.MAX_LEVELS)()
        self.keep, self.B, off = [], None, 0
        self.dtype, self.C = levels[0][1].dtype, levels[0][1].shape[-1]
        for i, (where, fvec, k) in enumerate(levels):
            sr = _as_search(where)
            self.keep.append((sr, fvec))
            self.B = sr.B if self.B is None else self.B
            if sr.B != self.B or fvec.dtype != self.dtype or not fvec.is_contiguous() or fvec.shape[-1] != self.C:
                raise MvtError("CorrPlan: levels must cover the same frames; fvec contiguous [B,N,C] of one dtype")
            self.lv[i] = CorrLevel(sr.c, fvec.data_ptr(), int(k), off)
            off += int(k) * 4
        self.width = off
        self.tiles = all(sr.c.tile_z for sr, _ in self.keep)
        self.fd = _feat_dtype(levels[0][1])
        self._peers = {}

    def peers_at(self, out_peers: "_lib.Peers", nbytes: int):
        key = (out_peers.ptr[0], out_peers.mc, nbytes)
        q = self._peers.get(key)
        if q is None:
            q = self._peers[key] = offset_peers(out_peers, nbytes)
        return q


def knn_corr_shared(levels, targets: torch.Tensor, coords: torch.Tensor, frame0: int, frame_step: int,
                    out_peers: "_lib.Peers", width: Optional[int] = None, defer="auto") -> None:
    """The fused kernel for THIS rank's frames of a window, storing into every rank's result block (mvt_knn_corr_shared).
    levels: (where, fvec, k) over the clouds of the B local frames -- or a CorrPlan made from them --, which are frames
    frame0, frame0 + frame_step, ... of targets [S, M, C] / coords [S, M, 3] (replicated, contiguous).  out_peers: the
    [S, M, width] fp32 block of every rank (its multicast address or the peer mappings; width = n_levels * k * 4 unless given)."""
    plan = levels if isinstance(levels, CorrPlan) else CorrPlan(levels)
    S, M, Cc = targets.shape
    if not (targets.is_cuda and coords.is_cuda and targets.is_contiguous() and coords.is_contiguous()) or coords.dtype != torch.float32:
        raise MvtError("knn_corr_shared: targets / coords must be contiguous CUDA tensors (coords fp32)")
    if targets.dtype != plan.dtype or Cc != plan.C:
        raise MvtError("knn_corr_shared: targets do not match the clouds' features")
    n, B = plan.n, plan.B
    width = plan.width if width is None else int(width)
    if B == 0:
        return
    if frame0 < 0 or frame_step < 1 or frame0 + (B - 1) * frame_step >= S or width < plan.width:
        raise MvtError(f"knn_corr_shared: frames {frame0} + i*{frame_step}, i < {B}, do not fit the {S}-frame window")
    ws, mode, pol = None, 0, None
    if defer and plan.tiles and not os.environ.get("MVT_NO_DEFER"):
        ws = _corr_workspace(targets.device, n, B, M)
        mode = 1
        if defer == "auto":
            pol = hard_query_policy(targets.device)
            mode = pol.mode()
    out = plan.peers_at(out_peers, frame0 * M * width * 4)
    check(lib().mvt_knn_corr_shared(plan.lv, n, B, M, Cc, plan.fd, targets.data_ptr() + frame0 * M * Cc * targets.element_size(),
                                    coords.data_ptr() + frame0 * M * 12, frame_step * M, C.byref(out), width,
                                    frame_step * M * width, None if ws is None else ws.data_ptr(), mode,
                                    torch.cuda.current_stream().cuda_stream), "mvt_knn_corr_shared")
    if pol is not None:
        pol.record(ws, n * B * M, mode)


def knn_corr_backward(fvec: torch.Tensor, targets: torch.Tensor, idx: torch.Tensor, grad_out: torch.Tensor, groups: int = 1,
                      grad_targets: Optional[torch.Tensor] = None, grad_fvec: Optional[torch.Tensor] = None):
    """Row f2: backward of one level of corr_sample (mvtracker.py:827-835) given the forward pass's neighbour lists.
    fvec [B,N,C], targets [B,M,C] (same dtype), idx [B,M,k] int32, grad_out [B,M,k,D] fp32 (any strides with a unit
    last one, e.g. the level's slice of the fused block's gradient) -> (grad_targets [B,M,C], grad_fvec [B,N,C]) fp32.
    Passing grad_targets / grad_fvec accumulates into them (further levels of the same targets / a shared cloud)."""
    _need_cuda(fvec, "fvec"); _need_cuda(targets, "targets"); _need_cuda(grad_out, "grad_out")
    B, N, C = fvec.shape
    _, M, k, D = grad_out.shape
    assert targets.shape == (B, M, C) and idx.shape == (B, M, k) and idx.dtype == torch.int32
    if fvec.dtype != targets.dtype:
        raise MvtError(f"targets dtype {targets.dtype} != fvec dtype {fvec.dtype}")
    fv, tg, ix = fvec.contiguous(), targets.contiguous(), idx.contiguous()
    go = grad_out if grad_out.dtype == torch.float32 else grad_out.float()
    if go.stride(3) != 1 or go.stride(2) != D:
        go = go.contiguous()
    acc = grad_targets is not None
    if grad_targets is None:
        grad_targets = torch.empty(B, M, C, device=fv.device, dtype=torch.float32)
    if grad_fvec is None:
        grad_fvec = torch.zeros(B, N, C, device=fv.device, dtype=torch.float32)
    for t in (grad_targets, grad_fvec):
        assert t.dtype == torch.float32 and t.is_contiguous()
    check(lib().mvt_knn_corr_backward(_ptr(fv), _ptr(tg), _feat_dtype(fv), _ptr(ix), _ptr(go), go.stride(1), go.stride(0), B, N, M, C,
                                      int(groups), k, D, _ptr(grad_targets), int(acc), _ptr(grad_fvec), _stream()),
          "mvt_knn_corr_backward")
    return grad_targets, grad_fvec


def cloud_build_backward(grad_fvec: torch.Tensor, V: int, H: int, W: int, level: int,
                         grad_fmaps: Optional[torch.Tensor] = None, accumulate: Optional[bool] = None) -> torch.Tensor:
    """Row f2: feature branch of init_pointcloud_from_rgbd backwards (model_utils.py:439-441, 475-476).
    grad_fvec [S, V*(H>>level)*(W>>level), C] -> grad_fmaps [V,S,C,H,W] fp32 (accumulated into when given)."""
    _need_cuda(grad_fvec, "grad_fvec")
    S, P, C = grad_fvec.shape
    assert P == V * (H >> level) * (W >> level), "grad_fvec does not match the level's cloud size"
    g = _f32c(grad_fvec)
    acc = (grad_fmaps is not None) if accumulate is None else bool(accumulate)
    if grad_fmaps is None:
        assert not acc
        grad_fmaps = torch.empty(V, S, C, H, W, device=g.device, dtype=torch.float32)
    assert grad_fmaps.dtype == torch.float32 and grad_fmaps.is_contiguous() and grad_fmaps.shape == (V, S, C, H, W)
    check(lib().mvt_cloud_build_backward(_ptr(g), V, S, C, H, W, int(level), _ptr(grad_fmaps), int(acc), _stream()),
          "mvt_cloud_build_backward")
    return grad_fmaps


def lift_pixels(depth: torch.Tensor, cam_record: torch.Tensor, pixels_xy: torch.Tensor, t: Optional[float] = None) -> torch.Tensor:
    """Row f3: depth [H,W] fp32, cam_record [CAM_STRIDE] (camera_prep), pixels_xy [n,2] -> world points [n,3], or
    [n,4] = (t, x, y, z) when t is given (evaluation_predictor_3dpt.py:99-116)."""
    _need_cuda(depth, "depth"); _need_cuda(pixels_xy, "pixels_xy"); _need_cuda(cam_record, "cam_record")
    H, W = depth.shape
    d, px, cm = _f32c(depth), _f32c(pixels_xy.reshape(-1, 2)), _f32c(cam_record.reshape(-1))
    assert cm.numel() == _lib.CAM_STRIDE
    out = torch.empty(px.shape[0], 3 if t is None else 4, device=d.device, dtype=torch.float32)
    check(lib().mvt_lift_pixels(_ptr(d), H, W, _ptr(cm), _ptr(px), px.shape[0], int(t is not None), float(t or 0.0), _ptr(out),
                                _stream()), "mvt_lift_pixels")
    return out


def track_input(coords: torch.Tensor, ffeats: torch.Tensor, track_mask: torch.Tensor, vis: torch.Tensor,
                pos_embed: Optional[torch.Tensor], time_embed: Optional[torch.Tensor], flow_embed_dim: int = 64,
                corr_width: int = 256, out: Optional[torch.Tensor] = None, corr_in_out: bool = False) -> torch.Tensor:
    """Row f1: the update transformer's input x [N, S, D] (mvtracker.py:374-386), D = 3E+3 + corr_width + C + 2.
    coords [S,N,3], ffeats [S,N,C], track_mask / vis [S,N] (or [S,N,1]), pos_embed [N,D] (or [N,1,D]), time_embed [S,D]
    (or [1,S,D]).  corr_in_out: `out` already holds the correlation block in columns [3E+3, 3E+3+corr_width) (written by
    knn_corr(out=corr_columns(out))) and gets the embeddings added there in place; otherwise those columns are untouched."""
    _need_cuda(coords, "coords"); _need_cuda(ffeats, "ffeats")
    S, N, C = ffeats.shape
    E = int(flow_embed_dim)
    D = 3 * E + 3 + int(corr_width) + C + 2
    cq, ff = _f32c(coords.reshape(S, N, 3)), ffeats.contiguous()
    tm, vs = _f32c(track_mask.reshape(S, N)), _f32c(vis.reshape(S, N))
    pe = _f32c(pos_embed.reshape(N, D)) if pos_embed is not None else None
    te = _f32c(time_embed.reshape(S, D)) if time_embed is not None else None
    if out is None:
        assert not corr_in_out
        out = torch.empty(N, S, D, device=ff.device, dtype=torch.float32)
    assert out.dtype == torch.float32 and out.shape == (N, S, D) and out.stride(2) == 1
    check(lib().mvt_track_input(_ptr(cq), _ptr(ff), _feat_dtype(ff), _ptr(tm), _ptr(vs), _ptr(pe), _ptr(te), S, N, C, E,
                                int(corr_width), _ptr(out), out.stride(0), out.stride(1), int(bool(corr_in_out)), _stream()),
          "mvt_track_input")
    return out


def corr_columns(x: torch.Tensor, flow_embed_dim: int = 64, corr_width: int = 256) -> torch.Tensor:
    """The [S, N, corr_width] view of a [N, S, D] transformer-input buffer that knn_corr(out=...) writes into."""
    c0 = 3 * int(flow_embed_dim) + 3
    return x.permute(1, 0, 2)[:, :, c0:c0 + corr_width]


def knn_gather(where, fvec: torch.Tensor, query: torch.Tensor, return_idx=False):
    """k=1: feature row of the nearest cloud point per query.  fvec [B,N,C] (or the frame's [1,N,C]), query [B,M,3]."""
    _need_cuda(query, "query"); _need_cuda(fvec, "fvec")
    sr = _as_search(where, tiles=True)
    q = _f32c(query)
    fv = fvec.contiguous()
    B, M, _ = q.shape
    if B != sr.B or fv.shape[0] != B or fv.shape[1] != sr.N:
        raise MvtError(f"shape mismatch: query batch {B}, search batch {sr.B}, fvec {tuple(fv.shape)}")
    out = torch.empty(B, M, fv.shape[-1], device=q.device, dtype=fv.dtype)
    idx = torch.empty(B, M, device=q.device, dtype=torch.int32) if return_idx else None
    check(lib().mvt_knn_gather(C.byref(sr.c), _ptr(fv), _feat_dtype(fv), B, fv.shape[-1], _ptr(q), M, _ptr(out), _ptr(idx),
                               _stream()), "mvt_knn_gather")
    return (out, idx) if return_idx else out


# ---------------------------------------------------------------------------------------------------------------------
# view-sharded path (one rank's halves of corr_sample; mvtracker_b200/dist.py adds the exchanges)
# ---------------------------------------------------------------------------------------------------------------------
EMPTY_KEY = 0x7F8000007FFFFFFF      # (+inf, no index): a missing neighbour in a packed key list


def knn_keys(searches: Sequence, index_offsets: Sequence[int], query: torch.Tensor, k: int, n_shards: int = 1,
             out: Optional[torch.Tensor] = None, caps: Optional[torch.Tensor] = None, budget: int = 0,
             exact_out: Optional[torch.Tensor] = None, skip: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Exact top-k of every query among this rank's points, for all levels in one launch.
    -> keys [n_shards, L, B, Ms, k] int64 storage of packed (bits(d2) << 32 | global index), ascending; Ms = ceil(M / n_shards).
    caps [n_shards, L, B, Ms] int64: per-query bound keys; budget / exact_out / skip (uint8, same layout): see mvt_knn_keys."""
    _need_cuda(query, "query")
    q = _f32c(query)
    B, M, _ = q.shape
    L = len(searches)
    srs = [_as_search(s, tiles=True) for s in searches]
    arr = (_CSearch * L)(*[s.c for s in srs])
    offs = (C.c_longlong * L)(*[int(o) for o in index_offsets])
    Ms = (M + n_shards - 1) // n_shards
    if out is None:
        out = torch.empty(n_shards, L, B, Ms, k, device=q.device, dtype=torch.int64)
    assert out.is_contiguous() and out.numel() == n_shards * L * B * Ms * k and out.dtype == torch.int64
    if caps is not None:
        assert caps.dtype == torch.int64 and caps.is_contiguous() and caps.numel() == n_shards * L * B * Ms
    for t in (exact_out, skip):
        if t is not None:
            assert t.dtype == torch.uint8 and t.is_contiguous() and t.numel() == n_shards * L * B * Ms
    check(lib().mvt_knn_keys(arr, offs, L, B, _ptr(q), M, int(k), int(n_shards), _ptr(caps), int(budget), _ptr(exact_out),
                             _ptr(skip), _ptr(out), _stream()), "mvt_knn_keys")
    return out


def topk_merge(keys: torch.Tensor, out: Optional[torch.Tensor] = None, keys_out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """keys [n_lists, Q, k] packed sorted lists -> idx [Q, k] int32 (global indices, -1 = missing) of the merged top-k.
    keys_out [Q, k] int64: receive the merged keys instead (then no index tensor is made and keys_out is returned)."""
    _need_cuda(keys, "keys")
    assert keys.dtype == torch.int64 and keys.is_contiguous() and keys.dim() == 3
    n_lists, Q, k = keys.shape
    if keys_out is not None:
        assert keys_out.dtype == torch.int64 and keys_out.is_contiguous() and keys_out.numel() == Q * k
        check(lib().mvt_topk_merge(_ptr(keys), n_lists, Q, k, None, _ptr(keys_out), _stream()), "mvt_topk_merge")
        return keys_out
    if out is None:
        out = torch.empty(Q, k, device=keys.device, dtype=torch.int32)
    check(lib().mvt_topk_merge(_ptr(keys), n_lists, Q, k, _ptr(out), None, _stream()), "mvt_topk_merge")
    return out


def corr_from_idx(owned: Sequence, idx: torch.Tensor, targets: torch.Tensor, coords: torch.Tensor, k: int,
                  out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """owned: per level (xyz4 [B,Nloc,4], fvec [B,Nloc,C], lo, hi); idx [n_shards, L, B, Ms, k] int32 global indices.
    -> out [B, M, L*k*4] fp32 with the records of the neighbours owned here, zeros elsewhere."""
    _need_cuda(targets, "targets"); _need_cuda(idx, "idx")
    B, M, Cc = targets.shape
    tg, cq = targets.contiguous(), _f32c(coords)
    L = len(owned)
    n_shards = idx.shape[0]
    assert idx.dtype == torch.int32 and idx.is_contiguous() and idx.shape[1] == L and idx.shape[2] == B and idx.shape[4] == k
    lv = (_lib.OwnedLevel * L)()
    for i, (xyz4, fvec, lo, hi) in enumerate(owned):
        if fvec.dtype != tg.dtype:
            raise MvtError(f"targets dtype {tg.dtype} != fvec dtype {fvec.dtype}")
        assert xyz4.is_contiguous() and fvec.is_contiguous() and xyz4.shape[1] == hi - lo
        lv[i] = _lib.OwnedLevel(xyz4.data_ptr(), fvec.data_ptr(), int(lo), int(hi), i * k * 4)
    if out is None:
        out = torch.empty(B, M, L * k * 4, device=tg.device, dtype=torch.float32)
    assert out.is_contiguous()
    check(lib().mvt_corr_from_idx(lv, L, B, M, Cc, _feat_dtype(tg), int(k), int(n_shards), _ptr(idx), _ptr(tg), _ptr(cq),
                                  _ptr(out), out.stride(1), _stream()), "mvt_corr_from_idx")
    return out


def full_cloud(level: int, xyz4: torch.Tensor, cam: torch.Tensor, V: int, H: int, W: int, stride: float) -> Cloud:
    """Search-only Cloud over replicated raster records [S, V*H*W, 4] (no features): what mvtracker_b200.dist searches
    in "tracks" mode."""
    return Cloud(level, None, None, None, None, xyz4, cam, V, H, W, stride)      # tile_z: ops.tile_depth_build([...])


def knn_idx_shared(searches: Sequence, query: torch.Tensor, k: int, shard: int, n_shards: int, peers) -> None:
    """Exact kNN of this rank's track shard (query [B, Ms, 3]); index lists stored into idx_all of every rank (peers)."""
    _need_cuda(query, "query")
    q = _f32c(query)
    B, Ms, _ = q.shape
    srs = [_as_search(s, tiles=True) for s in searches]
    arr = (_CSearch * len(srs))(*[s.c for s in srs])
    check(lib().mvt_knn_idx_shared(arr, len(srs), B, _ptr(q), Ms, int(k), int(shard), int(n_shards), C.byref(peers), _stream()),
          "mvt_knn_idx_shared")


def corr_from_idx_shared(owned: Sequence, idx: torch.Tensor, targets: torch.Tensor, coords: torch.Tensor, k: int, shard: int,
                         peers, row_stride: int, replicas: Optional[Sequence[torch.Tensor]] = None) -> None:
    """corr_from_idx whose records are stored into the [B, M, L*k*4] block of every rank (peers).
    replicas: per level, this rank's copy of every view's records [B, N_all, 4]; then only correlation values cross
    NVLink and each rank fills in the offsets itself."""
    _need_cuda(targets, "targets"); _need_cuda(idx, "idx")
    B, M, Cc = targets.shape
    tg, cq = targets.contiguous(), _f32c(coords)
    L = len(owned)
    assert idx.dtype == torch.int32 and idx.is_contiguous() and idx.shape[1] == L and idx.shape[2] == B and idx.shape[4] == k
    lv = (_lib.OwnedLevel * L)()
    for i, (xyz4, fvec, lo, hi) in enumerate(owned):
        if fvec.dtype != tg.dtype:
            raise MvtError(f"targets dtype {tg.dtype} != fvec dtype {fvec.dtype}")
        lv[i] = _lib.OwnedLevel(xyz4.data_ptr(), fvec.data_ptr(), int(lo), int(hi), i * k * 4)
    rp = na = None
    if replicas is not None:
        assert len(replicas) == L and all(r.is_contiguous() and r.dtype == torch.float32 and r.shape[0] == B for r in replicas)
        rp = (C.c_void_p * L)(*[r.data_ptr() for r in replicas])
        na = (C.c_longlong * L)(*[r.shape[1] for r in replicas])
    check(lib().mvt_corr_from_idx_shared(lv, L, B, M, Cc, _feat_dtype(tg), int(k), int(idx.shape[0]), int(shard), _ptr(idx),
                                         _ptr(tg), _ptr(cq), C.byref(peers), int(row_stride), rp, na, _stream()),
          "mvt_corr_from_idx_shared")


# ---------------------------------------------------------------------------------------------------------------------
# one kernel per refinement iteration across ranks (mvt_shard_iter)
# ---------------------------------------------------------------------------------------------------------------------
def plain_peers(t: torch.Tensor) -> "_lib.Peers":
    """A one-rank mvt_peers over an ordinary tensor (no multicast): what a single process passes where the multi-GPU
    path passes symmetric buffers."""
    p = _lib.Peers()
    p.n, p.mc = 1, None
    p.ptr[0] = t.data_ptr()
    return p


class ShardState:
    """The buffers of mvt_shard_iter for one rank: inbox [n_ranks, cap] int64, ctrl [16] int64, two alternating result
    blocks [B, M, L*k*4] fp32 (each given as (tensor, peers)), private counters, the epoch.  mvtracker_b200.dist allocates the
    first three in symmetric memory; with n_ranks == 1 plain tensors do (tests, single-GPU use)."""

    def __init__(self, rank: int, n_ranks: int, view_first: Sequence[int], inbox, ctrl, results, cap: int, device):
        self.rank, self.n, self.view_first = int(rank), int(n_ranks), [int(v) for v in view_first]
        self.inbox, self.ctrl, self.results, self.cap = inbox, ctrl, results, int(cap)
        B, M, _ = results[0][0].shape
        # private counters + the deferred-query list (sized for MAX_LEVELS levels of this rank's track shard)
        self.local = torch.zeros(int(lib().mvt_shard_local_bytes(_lib.MAX_LEVELS, B, M, self.n)), device=device, dtype=torch.uint8)
        self.epoch, self.flip = 0, 1

    @classmethod
    def single(cls, V: int, n_levels: int, B: int, M: int, k: int, device) -> "ShardState":
        cap = n_levels * B * M * k
        inbox = torch.zeros(1, cap, device=device, dtype=torch.int64)
        ctrl = torch.zeros(2 * _lib.Peers.ptr.size // C.sizeof(C.c_void_p), device=device, dtype=torch.int64)
        res = [torch.empty(B, M, n_levels * k * 4, device=device, dtype=torch.float32) for _ in range(2)]
        return cls(0, 1, [0, V], (inbox, plain_peers(inbox)), (ctrl, plain_peers(ctrl)), [(r, plain_peers(r)) for r in res], cap, device)

    def error(self) -> bool:
        """A wait inside the kernel timed out (a peer never arrived): the last results are invalid."""
        return bool(self.local[12:16].view(torch.int32).item())


def shard_iter(full: Sequence[Cloud], local: Sequence[Cloud], targets: torch.Tensor, coords: torch.Tensor, k: int,
               st: ShardState) -> torch.Tensor:
    """One refinement iteration's correlation block across ranks in ONE kernel launch per rank (mvt_shard_iter).
    full: per level, the search-only Cloud over the replicated records of all views (tile tables built);
    local: per level, this rank's Cloud (xyz4 + fvec of its views); targets [B,M,C] / coords [B,M,3]: all tracks.
    -> the result block [B, M, L*k*4] of this rank (complete when the kernel has finished; alternates between two buffers)."""
    _need_cuda(targets, "targets"); _need_cuda(coords, "coords")
    B, M, Cc = targets.shape
    tg, cq = targets.contiguous(), _f32c(coords)
    L = len(full)
    lv = (_lib.ShardLevel * L)()
    keep = []
    for i, (f, c) in enumerate(zip(full, local)):
        sr = Search.from_cloud(f, tiles=True)
        keep.append(sr)
        if c.fvec.dtype != tg.dtype:
            raise MvtError(f"targets dtype {tg.dtype} != fvec dtype {c.fvec.dtype}")
        lv[i] = _lib.ShardLevel(sr.c, c.xyz4.data_ptr(), c.fvec.data_ptr(), c.xyz4.shape[1], f.H * f.W)
    st.flip = 1 - st.flip
    st.epoch += 1
    res, res_peers = st.results[st.flip]
    ctx = _lib.ShardCtx()
    ctx.inbox, ctx.ctrl, ctx.result = st.inbox[1], st.ctrl[1], res_peers
    ctx.local, ctx.inbox_cap, ctx.rank, ctx.n_ranks = st.local.data_ptr(), st.cap, st.rank, st.n
    for r in range(9):
        ctx.view_first[r] = st.view_first[min(r, st.n)]
    check(lib().mvt_shard_iter(lv, L, B, M, Cc, _feat_dtype(tg), int(k), _ptr(tg), _ptr(cq), C.byref(ctx), res.stride(1),
                               st.epoch, _stream()), "mvt_shard_iter")
    return res
