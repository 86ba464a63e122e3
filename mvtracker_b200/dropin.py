"""Host-side mirror of the reference's seam for the hot path: same names, arguments, return shapes and assertion
behaviour as the three module-level objects that `MVTracker.forward` resolves at call time
(reference mvtracker/models/core/mvtracker/mvtracker.py:20, :87/:90, :312; called at :298, :608, :637, :816):

    init_pointcloud_from_rgbd   mvtracker/models/core/model_utils.py:420-482
    knn                         mvtracker/models/core/mvtracker/mvtracker.py:26-90
    PointcloudCorrBlock         mvtracker/models/core/mvtracker/mvtracker.py:769-846

`install()` rebinds them, so `MVTracker.forward(rgbs, depths, query_points, intrs, extrs)` is unchanged.
CUDA tensors only; CPU tensors raise (the reference itself is the CPU path).

Two memoisations keep the reference's call pattern (one call per pyramid level) from costing a launch per level:
  * the first init_pointcloud_from_rgbd(level=l) call on a given (fmaps, depths, intrs, extrs) builds levels 0..3 in one
    kernel pass; calls for the other levels return the cached tensors.  Entries hold references to their inputs and
    are keyed on (data_ptr, _version), so an in-place edit or a new tensor is a miss.
  * the first corr_sample of a refinement iteration runs all sibling levels in ONE fused launch; the sibling blocks'
    corr_sample calls with the same (targets, coords) return their column slice.
Set `settings.pyramid_cache = False` / `settings.fuse_levels = False` for the plain one-launch-per-call behaviour.
"""
import weakref
from types import SimpleNamespace
from typing import Callable, Optional

import torch

from . import ops
from ._lib import MAX_LEVELS, MvtError

settings = SimpleNamespace(
    pyramid_cache=True,
    fuse_levels=True,
    # 'reference': bit-compatible with mvtracker.py:820-829 (kNN on the compacted valid cloud, gather with those
    # indices from the uncompacted arrays).  'intended': invalid points are skipped and the true neighbours gathered.
    filter_invalid_mode="reference",
    cell_scale=0.0,
    # 'raster': clouds that came from init_pointcloud_from_rgbd are searched directly in their depth rasters (no index
    # to build); 'grid': always build the hashed grid.  Clouds of unknown origin always use the grid.
    search="raster",
    # Per-frame cloud cache across the overlapping windows of one sequence (hotpath.WindowSequence / ops.CloudArena): the
    # clouds of the S/2 frames a window shares with the previous one are kept instead of rebuilt.  Needs the provenance that
    # the smart_cat wrapper of install() records (a window's tensors are `cat(previous[:, :, S/2:], new)`,
    # mvtracker.py:549-575) -- it is proven, never guessed; anything unproven is rebuilt.  The tensors returned for such a
    # window are views of the arena: valid until a few windows later, which is all the model needs (mvtracker.py:296-376).
    frame_cache=False,
    window=12,                  # the model's sliding_window_len: calls with this many frames get the full pyramid
    arena_slots=None,           # frames of arena per sequence shape (default: up to 4 windows, a third of the free memory)
)


def _version(t: torch.Tensor):
    """In-place edit counter, or None for inference tensors (torch.inference_mode(): 'Inference tensors do not track
    version counter')."""
    try:
        return t._version
    except RuntimeError:
        return None


def _key(t: torch.Tensor):
    return (t.data_ptr(), _version(t), tuple(t.shape), tuple(t.stride()), t.dtype)


def _untouched(t: torch.Tensor) -> bool:
    """One of OUR output tensors has not been written since it was made.  Inference tensors have no counter; they are
    immutable outside inference mode and the model never edits the clouds (mvtracker.py:296-376), so they pass."""
    v = _version(t)
    return v is None or v == 0


class _Pyramid:
    """All levels built from one (fmaps, depths, intrs, extrs) window; also the sibling group of its corr blocks."""

    def __init__(self, key, inputs, clouds):
        self.key = key
        self.inputs = inputs            # keeps the storages alive so data_ptr cannot be recycled while cached
        self.clouds = clouds            # level -> ops.Cloud
        self.blocks = weakref.WeakValueDictionary()   # level -> PointcloudCorrBlock registered on these clouds
        self.corr_cache = None          # (key, {level: tensor})
        self.served = set()             # levels already handed out (see init_pointcloud_from_rgbd)
        self.versioned = all(k[1] is not None for k in key[:4])
        self.seq = None                 # frame cache: the WindowSequence (arena) these clouds are views of


_pyramid_cache = []     # most recent first, at most 2 entries (feature-init call + forward_iteration call per window)
_arenas = {}            # (V, H, W, C, dtype, device, stride, mode) -> hotpath.WindowSequence, kept across forward calls
_ref_smart_cat = None   # the reference's own smart_cat (model_utils.py), wrapped by install()


def clear_caches(arenas: bool = False):
    _pyramid_cache.clear()
    _pending_blocks.clear()
    for seq in _arenas.values():
        seq.slot0 = None
    if arenas:
        _arenas.clear()


# ----------------------------------------------------------------------------------------------------------------------
# frame cache: provenance of a rolled window
# ----------------------------------------------------------------------------------------------------------------------
def _is_tail_of(t1: torch.Tensor, base: torch.Tensor, hop: int) -> bool:
    """t1 is exactly base[:, :, hop:] (same storage, same strides) of a [B, V, S, ...] tensor."""
    if base is None or t1.dim() != base.dim() or t1.dim() < 3 or t1.dtype != base.dtype:
        return False
    shape = list(base.shape)
    shape[2] -= hop
    # (`base` is alive, so its storage is: an equal storage pointer is the same storage, not a recycled address)
    return (t1.untyped_storage().data_ptr() == base.untyped_storage().data_ptr() and list(t1.shape) == shape
            and t1.stride() == base.stride() and t1.storage_offset() == base.storage_offset() + hop * base.stride(2))


def smart_cat(tensor1, tensor2, dim):
    """The reference's smart_cat (model_utils.py) with the provenance of a rolled window attached to its result: when
    tensor1 is the last S - hop frames of the tensor the CURRENT arena window was built from, the result's first S - hop
    frames are frames whose clouds are already in the arena.  The result itself is the reference's, untouched."""
    out = _ref_smart_cat(tensor1, tensor2, dim)
    try:
        if settings.frame_cache and tensor1 is not None and dim == 2 and out is not tensor2 and _pyramid_cache:
            cur = _pyramid_cache[0]
            if cur.seq is not None and cur.seq.current is cur:
                hop = out.shape[2] - tensor1.shape[2]
                for which, base in (("f", cur.fm), ("d", cur.dp)):
                    if base is not None and _version(base) == cur.versions[which] and _is_tail_of(tensor1, base, hop):
                        out._mvt_roll = (weakref.ref(cur), which, hop)
    except Exception:       # provenance is an optimisation: never let it break the model's own cat
        pass
    return out


def _same_camera_tail(new: torch.Tensor, old: torch.Tensor, old_version, hop: int) -> bool:
    """new[:, :, :S-hop] is old[:, :, hop:] -- both are slices of one untouched tensor (mvtracker.py:547-548)."""
    try:
        return (new.shape == old.shape and new.stride() == old.stride() and new.dtype == old.dtype
                and new.untyped_storage().data_ptr() == old.untyped_storage().data_ptr()
                and new.storage_offset() == old.storage_offset() + hop * old.stride(2)
                and _version(new) is not None and _version(new) == old_version)
    except Exception:
        return False


def _arena_for(fmaps, stride, mode):
    from .hotpath import WindowSequence
    _, V, S, C, H, W = fmaps.shape
    key = (V, H, W, C, fmaps.dtype, fmaps.device, float(stride), mode, S)
    seq = _arenas.get(key)
    if seq is None:
        seq = WindowSequence(V, H, W, C, fmaps.dtype, fmaps.device, S=S, hop=S // 2, stride=stride, slots=settings.arena_slots,
                             want_valid=True, depth_interp_mode=mode)
        seq.current = None
        _arenas[key] = seq
    return seq


def _build_in_arena(fmaps, depths, intrs, extrs, stride, mode, key):
    """Window of `settings.window` frames, B == 1, raster search: through the per-frame cache.  -> _Pyramid"""
    S = fmaps.shape[2]
    seq = _arena_for(fmaps, stride, mode)
    rf, rd = getattr(fmaps, "_mvt_roll", None), getattr(depths, "_mvt_roll", None)
    prev = rf[0]() if rf is not None else None
    hop = rf[2] if rf is not None else 0
    rolled = (prev is not None and rd is not None and rd[0]() is prev and rf[1] == "f" and rd[1] == "d" and rd[2] == hop
              and prev.seq is seq and seq.current is prev and seq.slot0 is not None and hop == seq.hop
              and _version(fmaps) == 0 and _version(depths) == 0            # the cat results have not been written since
              and _same_camera_tail(intrs, prev.cams[0], prev.versions["K"], hop)
              and _same_camera_tail(extrs, prev.cams[1], prev.versions["E"], hop))
    if rolled:
        win = seq.push(fmaps[0][:, S - hop:], depths[0][:, S - hop:, 0], intrs[0], extrs[0])
    else:
        seq.slot0 = None                                                    # a new sequence (or nothing could be proven)
        win = seq.push(fmaps[0], depths[0][:, :, 0], intrs[0], extrs[0])
    pyr = _Pyramid(key, None, {c.level: c for c in win.clouds})
    # The window's input tensors are held until the next window replaces this entry: that is when the model lets go of them
    # too (its next window is cat(this[:, :, S/2:], new), mvtracker.py:549-575), so the peak memory is the reference's; they
    # make the provenance of the next window checkable (same storage, alive, unwritten).
    pyr.seq, pyr.fm, pyr.dp = seq, fmaps, depths
    pyr.cams = (intrs, extrs)
    pyr.versions = {"f": _version(fmaps), "d": _version(depths), "K": _version(intrs), "E": _version(extrs)}
    pyr.rolled = rolled
    seq.current = pyr
    return pyr


def init_pointcloud_from_rgbd(fmaps, depths, intrs, extrs, stride=4, level=0, depth_interp_mode="nearest",
                              return_validity_mask=False):
    B, V, S, C, H, W = fmaps.shape
    assert fmaps.shape == (B, V, S, C, H, W)
    assert depths.shape == (B, V, S, 1, H, W)
    assert intrs.shape == (B, V, S, 3, 3)
    assert extrs.shape == (B, V, S, 3, 4)
    if not fmaps.is_cuda:
        raise MvtError("mvtracker_b200.init_pointcloud_from_rgbd needs CUDA tensors (no CPU fallback)")
    if depth_interp_mode not in ("nearest", "avg"):
        if level == 0:
            depth_interp_mode = "nearest"      # callers pass "N/A" at level 0 (evaluator_3dpt.py:309-317)
        else:
            raise NotImplementedError
    if level >= MAX_LEVELS:
        raise MvtError(f"pyramid level {level} not supported (0..{MAX_LEVELS - 1})")
    out_dtype = fmaps.dtype
    key = (_key(fmaps), _key(depths), _key(intrs), _key(extrs), float(stride), depth_interp_mode)
    if fmaps.dtype not in (torch.float32, torch.bfloat16):
        # fp16 / fp64 feature maps (fp16: autocast on pre-Ampere devices, demo.py:833-837): the kernels take fp32 or bf16
        # features; such maps go through the fp32 kernels and the cloud's features are returned in the caller's dtype
        fmaps_cast = fmaps.float()
        for name in ("_mvt_roll",):
            if hasattr(fmaps, name):
                setattr(fmaps_cast, name, getattr(fmaps, name))
        fmaps = fmaps_cast
    pyr = None
    if settings.pyramid_cache:
        for e in _pyramid_cache:
            # inference tensors carry no version counter, so an in-place edit of the inputs cannot be seen: an entry built
            # from them serves every level ONCE (the four per-level calls of one window, mvtracker.py:297-306); a second
            # request for a level means a new window in the same buffers and is rebuilt
            if e.key == key and level in e.clouds and (e.versioned or level not in e.served):
                pyr = e
                break
    use_arena = (pyr is None and settings.frame_cache and settings.pyramid_cache and settings.search == "raster" and B == 1
                 and S == settings.window and S % 2 == 0 and (H >> (MAX_LEVELS - 1)) > 0 and (W >> (MAX_LEVELS - 1)) > 0
                 and all(k[1] is not None for k in key[:4]) and not (torch.is_grad_enabled() and fmaps.requires_grad))
    if use_arena:
        pyr = _build_in_arena(fmaps, depths, intrs, extrs, stride, depth_interp_mode, key)
        _pyramid_cache[:] = [pyr] + [e for e in _pyramid_cache if e.key != key and e.seq is not pyr.seq][:1]
    if pyr is None:
        if settings.pyramid_cache and (S == settings.window or not settings.frame_cache):
            levels = [l for l in range(MAX_LEVELS) if (H >> l) > 0 and (W >> l) > 0]
            if level not in levels:
                levels = [level]
        else:
            # (frame cache on: a call with another number of frames is the feature-init call on the new frames,
            # mvtracker.py:606-613, which only ever asks for level 0 -- build what is asked for)
            levels = [level]
        per_b = []
        for b in range(B):
            per_b.append(ops.cloud_build(fmaps[b], depths[b, :, :, 0], intrs[b], extrs[b], stride=stride, levels=levels,
                                         depth_interp_mode=depth_interp_mode, want_valid=True,
                                         want_tiles=(settings.search == "raster" and B == 1)))
        clouds = {}
        for i, l in enumerate(levels):
            if B == 1:
                clouds[l] = per_b[0][i]
            else:
                cs = [pb[i] for pb in per_b]
                clouds[l] = ops.Cloud(l, torch.cat([c.xyz for c in cs]), torch.cat([c.fvec for c in cs]),
                                      torch.cat([c.valid for c in cs]), torch.cat([c.spacing for c in cs]))
        pyr = _Pyramid(key, (fmaps, depths, intrs, extrs), clouds)
        if settings.pyramid_cache:
            _pyramid_cache[:] = [pyr] + [e for e in _pyramid_cache if e.key != key][:1]
    pyr.served.add(level)
    cloud = pyr.clouds[level]
    xyz, fvec = cloud.xyz, cloud.fvec
    if torch.is_grad_enabled() and fmaps.requires_grad:
        # training (row f2): the cloud's features are a function of fmaps (chained avg_pool2d + permute); depths, cameras
        # and therefore xyz carry no gradient in the reference either (they come from the data loader)
        fvec = _FeatureCloudFn.apply(fmaps, fvec, level)
    if fvec.dtype != out_dtype:
        fvec = fvec.to(out_dtype)
    xyz._mvt = (weakref.ref(pyr), level)   # lets PointcloudCorrBlock find the footprint hint and its sibling levels
    if return_validity_mask:
        return xyz, fvec, cloud.valid
    return xyz, fvec


class _FeatureCloudFn(torch.autograd.Function):
    """fvec of one pyramid level as a differentiable function of fmaps (model_utils.py:439-441, 475-476); the forward
    value is the cloud that mvt_cloud_build already produced, the backward is mvt_cloud_build_backward."""

    @staticmethod
    def forward(ctx, fmaps, fvec_built, level):
        ctx.level, ctx.shape, ctx.dtype = level, tuple(fmaps.shape), fmaps.dtype
        return fvec_built.view_as(fvec_built)

    @staticmethod
    def backward(ctx, grad_fvec):
        B, V, S, C, H, W = ctx.shape
        g = grad_fvec.reshape(B, S, -1, C)
        out = torch.stack([ops.cloud_build_backward(g[b], V, H, W, ctx.level) for b in range(B)])
        return out.to(ctx.dtype), None, None


class _CorrSampleFn(torch.autograd.Function):
    """corr_sample of ONE level with gradients w.r.t. targets and fvec (mvtracker.py:827-835); the neighbour search, the
    offsets and the positions are not differentiable (coords are detached in the reference, mvtracker.py:351)."""

    @staticmethod
    def forward(ctx, targets, fvec, block, coords):
        out, idx = ops.knn_corr([(block._search(), block._fvec_c.detach(), block.k)], targets.detach(), coords.detach(),
                                groups=block.groups, add_offset=block.corr_add_neighbor_offset,
                                add_xyz=block.corr_add_neighbor_xyz, return_idx=True)
        ctx.save_for_backward(targets, fvec, idx[0])
        ctx.groups = block.groups
        D = block.groups + 3 * bool(block.corr_add_neighbor_offset) + 3 * bool(block.corr_add_neighbor_xyz)
        return out.reshape(targets.shape[0], targets.shape[1], block.k, D)

    @staticmethod
    def backward(ctx, grad_out):
        targets, fvec, idx = ctx.saved_tensors
        gt, gf = ops.knn_corr_backward(fvec.detach(), targets.detach(), idx, grad_out, groups=ctx.groups)
        return gt.to(targets.dtype), gf.to(fvec.dtype), None, None


def knn(k: int, xyz_ref: torch.Tensor, xyz_query: torch.Tensor):
    """(B,N,3), (B,M,3) -> (dists (B,M,k) in xyz_ref.dtype, idx (B,M,k) int64), ascending distance."""
    if not xyz_ref.is_cuda:
        raise MvtError("mvtracker_b200.knn needs CUDA tensors (no CPU fallback)")
    B, N, _ = xyz_ref.shape
    assert xyz_query.shape[0] == B and xyz_query.shape[2] == 3
    where = _find_raster(xyz_ref) if settings.search == "raster" else None
    if where is None:
        where = ops.build_grids([(xyz_ref, None, None)], k_hint=k, cell_scale=settings.cell_scale)[0]
    dists, idx = ops.knn(where, xyz_query, k, idx_dtype=torch.int64)
    return dists.to(xyz_ref.dtype), idx


def _find_raster(xyz_ref: torch.Tensor):
    """If xyz_ref is (a run of whole frames of) a cloud produced by init_pointcloud_from_rgbd that is still cached --
    e.g. `rgbd_xyz[b, t].reshape(-1, 3)[None]` in the feature-init loop (mvtracker.py:633-637) -- return the raster
    search over exactly those frames; else None."""
    if xyz_ref.dtype != torch.float32 or not xyz_ref.is_contiguous():
        return None
    Bq, N, _ = xyz_ref.shape
    ptr = xyz_ref.data_ptr()
    for pyr in _pyramid_cache:
        for cloud in pyr.clouds.values():
            S, P = cloud.xyz.shape[0], cloud.xyz.shape[1]
            base = cloud.xyz.data_ptr()
            if cloud.xyz4 is None or P != N or not _untouched(cloud.xyz) or not _untouched(cloud.xyz4):
                continue
            off = ptr - base
            if off < 0 or off % (P * 12) != 0:
                continue
            f0 = off // (P * 12)
            if f0 + Bq <= S:
                return ops.Search.from_cloud(cloud, frame=int(f0), nframes=int(Bq), tiles=True)
    return None


_pending_blocks = weakref.WeakSet()   # blocks whose grids have not been built yet (built together on first use)


class PointcloudCorrBlock:
    def __init__(
            self,
            k: int,
            groups,
            xyz: torch.Tensor,
            fvec: torch.Tensor,
            corr_add_neighbor_offset: bool,
            corr_add_neighbor_xyz: bool,
            filter_invalid: bool = False,
            valid: Optional[torch.Tensor] = None,
            rerun_fmap_coloring_fn: Optional[Callable] = None,
    ):
        self.B, self.N, self.C = fvec.shape
        assert xyz.shape == (self.B, self.N, 3)
        assert fvec.shape == (self.B, self.N, self.C)
        assert k <= self.N, "k should be less than or equal to N"
        assert groups <= self.C, "number of correlation groups should not be larger than the number of channels"
        assert self.C % groups == 0, "number of channels must be divisible by the number of groups (for convenience)"
        assert not filter_invalid or valid is not None
        if not xyz.is_cuda:
            raise MvtError("mvtracker_b200.PointcloudCorrBlock needs CUDA tensors (no CPU fallback)")

        self.k = k
        self.groups = groups
        self.xyz = xyz
        self.fvec = fvec
        self.corr_add_neighbor_offset = corr_add_neighbor_offset
        self.corr_add_neighbor_xyz = corr_add_neighbor_xyz
        self.filter_invalid = filter_invalid
        self.valid = valid
        self.rerun_fmap_coloring_fn = rerun_fmap_coloring_fn

        self._xyz_c = xyz if (xyz.dtype == torch.float32 and xyz.is_contiguous()) else xyz.float().contiguous()
        self._grid = None
        self._pyr, self._level, self._spacing, self._cloud = None, None, None, None
        tag = getattr(xyz, "_mvt", None)
        if tag is not None and tag[0]() is not None and _untouched(xyz):
            self._pyr, self._level = tag[0](), tag[1]
            self._cloud = self._pyr.clouds[self._level]
            self._spacing = self._cloud.spacing
            self._pyr.blocks[self._level] = self
        if self._cloud is not None and self._cloud.fvec is not None and self._cloud.fvec.shape == fvec.shape and _untouched(fvec) \
                and (fvec.dtype == self._cloud.fvec.dtype or fvec.dtype not in (torch.float32, torch.bfloat16)):
            self._fvec_c = self._cloud.fvec          # the rows the kernels built (fp32 for fp16 / fp64 callers)
        else:
            self._fvec_c = fvec.contiguous() if fvec.dtype in (torch.float32, torch.bfloat16) else fvec.float().contiguous()
        self._raster = (settings.search == "raster" and self._cloud is not None and self._cloud.xyz4 is not None
                        and not (filter_invalid and settings.filter_invalid_mode == "reference"))
        if not self._raster:
            _pending_blocks.add(self)

    def _search(self):
        if self._raster:
            return ops.Search.from_cloud(self._cloud, skip_invalid=self.filter_invalid)
        self._ensure_grid()
        return ops.Search.from_grid(self._grid, self._xyz_c)

    # -- grids of all blocks created since the last build go into one mvt_grid_build call -------------------------
    def _ensure_grid(self):
        if self._grid is not None:
            return
        todo = [b for b in list(_pending_blocks) if b is not self and b._grid is None
                and b.filter_invalid == self.filter_invalid and b.k == self.k and b.xyz.device == self.xyz.device]
        todo = [self] + todo[:7]
        mode = 1 if (self.filter_invalid and settings.filter_invalid_mode == "reference") else 0
        jobs = [(b._xyz_c, b.valid if b.filter_invalid else None, b._spacing) for b in todo]
        grids = ops.build_grids(jobs, k_hint=self.k, index_mode=mode, cell_scale=settings.cell_scale)
        for b, g in zip(todo, grids):
            b._grid = g
            _pending_blocks.discard(b)

    def _siblings(self):
        """Levels of the same pyramid that can share one fused launch with this block."""
        if not settings.fuse_levels or self._pyr is None:
            return [self]
        sib = [b for _, b in sorted(self._pyr.blocks.items())
               if b.B == self.B and b.C == self.C and b.groups == self.groups and b.fvec.dtype == self.fvec.dtype
               and b.corr_add_neighbor_offset == self.corr_add_neighbor_offset
               and b.corr_add_neighbor_xyz == self.corr_add_neighbor_xyz and b.filter_invalid == self.filter_invalid
               and b._raster == self._raster]
        return sib if self in sib else [self]

    def corr_sample(
            self,
            targets: torch.Tensor,
            coords_world_xyz: torch.Tensor,
            save_debug_logs=False,
            debug_logs_path=".",
            debug_logs_prefix="corr",
            save_rerun_logs=False,
    ):
        _, M, _ = targets.shape
        assert targets.shape == (self.B, M, self.C)
        assert coords_world_xyz.shape == (self.B, M, 3)
        if save_debug_logs or save_rerun_logs:
            raise MvtError("debug / rerun logging lives in the reference implementation only")
        D = self.groups + 3 * bool(self.corr_add_neighbor_offset) + 3 * bool(self.corr_add_neighbor_xyz)
        out_dtype = torch.promote_types(targets.dtype, self.fvec.dtype)
        if self.corr_add_neighbor_offset or self.corr_add_neighbor_xyz:     # torch.cat promotion at mvtracker.py:842-846
            out_dtype = torch.promote_types(out_dtype, torch.promote_types(self.xyz.dtype, coords_world_xyz.dtype))
        ckey = (_key(targets), _key(coords_world_xyz))      # of the caller's tensors: the siblings are called with the same ones
        keep = (targets, coords_world_xyz)
        if targets.dtype != self._fvec_c.dtype:
            targets = targets.to(self._fvec_c.dtype)
        if torch.is_grad_enabled() and (targets.requires_grad or self.fvec.requires_grad):
            # training (row f2): one level per call, neighbour lists kept for mvt_knn_corr_backward
            return _CorrSampleFn.apply(targets, self.fvec, self, coords_world_xyz)

        sib = self._siblings()
        if len(sib) > 1:
            cache = self._pyr.corr_cache
            if cache is not None and cache[0] == ckey and id(self) in cache[1]:
                hit = cache[1].pop(id(self))
                return hit if hit.dtype == out_dtype else hit.to(out_dtype)
        levels = [(b._search(), b._fvec_c, b.k) for b in sib]
        out = ops.knn_corr(levels, targets, coords_world_xyz, groups=self.groups, add_offset=self.corr_add_neighbor_offset,
                           add_xyz=self.corr_add_neighbor_xyz)
        res, off = {}, 0
        for b in sib:
            w = b.k * D
            res[id(b)] = out[:, :, off:off + w].reshape(self.B, M, b.k, D)
            off += w
        mine = res.pop(id(self))
        if len(sib) > 1:
            # (targets, coords) are referenced by the key tensors' storages for the lifetime of this entry
            self._pyr.corr_cache = (ckey, res, keep, out_dtype)
        return mine if mine.dtype == out_dtype else mine.to(out_dtype)


def install(ref_module=None, frame_cache: bool = True, window: Optional[int] = 12):
    """Rebind the reference's three seam names to this implementation.  Returns a callable that restores them.
    frame_cache: also wrap the module's `smart_cat` (it keeps returning the reference's result) so that the clouds of the
    frames consecutive windows share are built once (settings.frame_cache); window: the model's sliding_window_len."""
    if ref_module is None:
        import importlib
        ref_module = importlib.import_module("mvtracker.models.core.mvtracker.mvtracker")
    global _ref_smart_cat
    names = ["init_pointcloud_from_rgbd", "knn", "PointcloudCorrBlock"]
    if frame_cache and hasattr(ref_module, "smart_cat"):
        names.append("smart_cat")
    saved = {n: getattr(ref_module, n) for n in names}
    saved_settings = (settings.frame_cache, settings.window)
    ref_module.init_pointcloud_from_rgbd = init_pointcloud_from_rgbd
    ref_module.knn = knn
    ref_module.PointcloudCorrBlock = PointcloudCorrBlock
    if "smart_cat" in saved:
        # a fourth, passive name: the wrapper returns the reference's own result and only records where it came from
        if saved["smart_cat"] is not smart_cat:
            _ref_smart_cat = saved["smart_cat"]
        ref_module.smart_cat = smart_cat
        settings.frame_cache = True
        if window is not None:
            settings.window = int(window)

    def uninstall():
        for n, v in saved.items():
            setattr(ref_module, n, v)
        settings.frame_cache, settings.window = saved_settings
        clear_caches()

    return uninstall
