"""Torch-facing wrappers over the C ABI (include/mvt.h).  PyTorch is plumbing only here: it owns device memory and the
current stream; every computation below runs in libmvt.so's sm_100a kernels.  CUDA tensors only -- there is no CPU path."""
from dataclasses import dataclass
from typing import List, Optional, Sequence

import torch

from . import _lib
from ._lib import C, CorrLevel, GridJob, LevelOut, MvtError, check, lib


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
    """intrs [...,3,3], extrs [...,3,4] -> cam [n,24] fp32 (Kinv | Einv rows), see mvt_camera_prep."""
    _need_cuda(intrs, "intrs"); _need_cuda(extrs, "extrs")
    K = intrs.reshape(-1, 3, 3).float().contiguous()
    E = extrs.reshape(-1, 3, 4).float().contiguous()
    assert K.shape[0] == E.shape[0]
    cam = torch.empty(K.shape[0], 24, device=K.device, dtype=torch.float32)
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


def cloud_build(fmaps, depths, intrs, extrs, stride=4, levels: Sequence[int] = (0,), depth_interp_mode="nearest",
                want_valid=False, want_fvec=True, cam=None) -> List[Cloud]:
    """fmaps [V,S,C,H,W], depths [V,S,H,W] (or [V,S,1,H,W]) -> one Cloud per requested level, built in ONE kernel pass."""
    _need_cuda(fmaps, "fmaps"); _need_cuda(depths, "depths")
    V, S, Cc, H, W = fmaps.shape
    depths = depths.reshape(V, S, H, W)
    if depth_interp_mode not in ("nearest", "avg"):
        raise NotImplementedError(depth_interp_mode)
    levels = sorted(set(int(l) for l in levels))
    if not levels or levels[0] < 0 or levels[-1] >= _lib.MAX_LEVELS:
        raise MvtError(f"levels {levels} outside 0..{_lib.MAX_LEVELS - 1}")
    fm = fmaps.contiguous()
    dp = depths.float().contiguous()
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
        xyz = torch.empty(S, P, 3, device=dev, dtype=torch.float32)
        fvec = torch.empty(S, P, Cc, device=dev, dtype=fm.dtype) if want_fvec else None
        valid = torch.empty(S, P, device=dev, dtype=torch.uint8) if want_valid else None
        spacing = torch.empty(S, 2, device=dev, dtype=torch.float32)
        outs[l] = LevelOut(xyz.data_ptr(), fvec.data_ptr() if want_fvec else None,
                           valid.data_ptr() if want_valid else None, spacing.data_ptr())
        mask |= 1 << l
        clouds.append(Cloud(l, xyz, fvec, valid.view(torch.bool) if want_valid else None, spacing))
    check(lib().mvt_cloud_build(_ptr(fm), _feat_dtype(fm), _ptr(dp), _ptr(cam), V, S, Cc, H, W, float(stride),
                                int(depth_interp_mode == "avg"), outs, mask, _stream()), "mvt_cloud_build")
    return clouds


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


def knn(grid: Grid, query: torch.Tensor, k: int, idx_dtype=torch.int64, want_dist=True):
    """-> (dists [B,M,k] fp32 ascending, idx [B,M,k]); canonical (d2, index) order; -1 / inf where fewer than k points."""
    _need_cuda(query, "query")
    q = query if (query.dtype == torch.float32 and query.is_contiguous()) else query.float().contiguous()
    B, M, _ = q.shape
    if B != grid.B:
        raise MvtError(f"query batch {B} != grid batch {grid.B}")
    idx = torch.empty(B, M, k, device=q.device, dtype=idx_dtype)
    dist = torch.empty(B, M, k, device=q.device, dtype=torch.float32) if want_dist else None
    check(lib().mvt_knn(C.c_void_p(grid.ptr), B, grid.N, _ptr(q), M, int(k), _ptr(idx), int(idx_dtype == torch.int64),
                        _ptr(dist), _stream()), "mvt_knn")
    return dist, idx


def knn_corr(levels: Sequence, targets: torch.Tensor, coords: torch.Tensor, groups=1, add_offset=True, add_xyz=False,
             out: Optional[torch.Tensor] = None, return_idx=False):
    """levels: sequence of (grid, xyz [B,N,3], fvec [B,N,C], k).  -> out [B, M, sum_l k_l*D] fp32 (level-major columns,
    the layout of torch.cat(fcorrs, -1) at mvtracker.py:374), D = groups + 3*add_offset + 3*add_xyz."""
    _need_cuda(targets, "targets"); _need_cuda(coords, "coords")
    B, M, Cc = targets.shape
    tg = targets.contiguous()
    cq = coords if (coords.dtype == torch.float32 and coords.is_contiguous()) else coords.float().contiguous()
    D = groups + 3 * bool(add_offset) + 3 * bool(add_xyz)
    n = len(levels)
    lv = (CorrLevel * _lib.MAX_LEVELS)()
    off = 0
    for i, (grid, xyz, fvec, k) in enumerate(levels):
        if fvec.dtype != tg.dtype:
            raise MvtError(f"targets dtype {tg.dtype} != fvec dtype {fvec.dtype}")
        if not (fvec.is_contiguous() and xyz.is_contiguous() and xyz.dtype == torch.float32):
            raise MvtError("xyz / fvec must be contiguous (xyz fp32)")
        lv[i] = CorrLevel(grid.ptr, xyz.data_ptr(), fvec.data_ptr(), grid.N, int(k), off)
        off += int(k) * D
    if out is None:
        out = torch.empty(B, M, off, device=tg.device, dtype=torch.float32)
    row_stride = out.stride(1)
    assert out.shape[0] == B and out.shape[1] == M and out.stride(2) == 1 and out.stride(0) == M * row_stride
    idx = None
    if return_idx:
        idx = torch.empty(n, B, M, int(levels[0][3]), device=tg.device, dtype=torch.int32)
    check(lib().mvt_knn_corr(lv, n, B, M, Cc, int(groups), _feat_dtype(tg), _ptr(tg), _ptr(cq), int(bool(add_offset)),
                             int(bool(add_xyz)), _ptr(out), row_stride, _ptr(idx), _stream()), "mvt_knn_corr")
    return (out, idx) if return_idx else out


def knn_gather(grid: Grid, fvec: torch.Tensor, query: torch.Tensor, return_idx=False):
    """k=1: feature row of the nearest cloud point per query.  fvec [B,N,C], query [B,M,3] -> [B,M,C]."""
    _need_cuda(query, "query"); _need_cuda(fvec, "fvec")
    q = query if (query.dtype == torch.float32 and query.is_contiguous()) else query.float().contiguous()
    fv = fvec.contiguous()
    B, M, _ = q.shape
    out = torch.empty(B, M, fv.shape[-1], device=q.device, dtype=fv.dtype)
    idx = torch.empty(B, M, device=q.device, dtype=torch.int32) if return_idx else None
    check(lib().mvt_knn_gather(C.c_void_p(grid.ptr), _ptr(fv), _feat_dtype(fv), B, grid.N, fv.shape[-1], _ptr(q), M,
                               _ptr(out), _ptr(idx), _stream()), "mvt_knn_gather")
    return (out, idx) if return_idx else out
