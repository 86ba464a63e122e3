"""torch.library registration of the hot path: `torch.ops.mvt.{cloud_build, knn, knn_corr, track_input}`.

SURVEY.md 8b names the seam underneath the three Python names as "torch extension ops mvt::cloud_build, mvt::knn,
mvt::knn_corr -> C-ABI libmvt functions".  These are those ops: opaque custom operators (torch.library.custom_op) whose
CUDA implementations call the same C ABI as mvtracker_b200.ops (include/mvt.h) and whose fake (meta) implementations give
shapes and dtypes, so that a surrounding model can be traced / compiled (torch.compile, torch.export, FakeTensorMode)
without a graph break at a ctypes call, and captured in a torch.cuda.CUDAGraph (every entry point only enqueues on the
current stream; tests/test_gpu_input.py::test_refinement_iteration_in_a_cuda_graph).  No CPU implementation is registered:
CPU tensors fail with the dispatcher's "no kernel for CPU" error -- there is no CPU fallback.

Import this module to register the ops (mvtracker_b200.dropin does not need them; they are the compile-friendly surface).
"""
from typing import List

import torch

from . import ops

_lib_def = torch.library.Library("mvt", "DEF")
_lib_def.define("cloud_build(Tensor fmaps, Tensor depths, Tensor intrs, Tensor extrs, float stride, int[] levels, "
                "bool depth_avg) -> Tensor[]")
_lib_def.define("knn(Tensor xyz4, Tensor cam, Tensor query, int k, int V, int H, int W, float stride) -> (Tensor, Tensor)")
_lib_def.define("knn_corr(Tensor[] xyz4, Tensor[] fvec, Tensor cam, Tensor targets, Tensor coords, int k, int V, int H, int W, "
                "float stride) -> Tensor")
_lib_def.define("track_input(Tensor coords, Tensor ffeats, Tensor track_mask, Tensor vis, Tensor pos_embed, Tensor time_embed, "
                "Tensor corr, int flow_embed_dim) -> Tensor")


def _cloud(level, xyz4, cam, V, H, W, stride, fvec=None):
    c = ops.Cloud(level, None, fvec, None, None, xyz4, cam, V, H >> level, W >> level, float(stride) * (1 << level))
    return c


# ---- cloud_build: -> per requested level [xyz, fvec, xyz4] (3 tensors per level, level-major) followed by cam ----------------
def _cloud_build_cuda(fmaps, depths, intrs, extrs, stride, levels, depth_avg):
    clouds = ops.cloud_build(fmaps, depths, intrs, extrs, stride=stride, levels=list(levels),
                             depth_interp_mode="avg" if depth_avg else "nearest", want_spacing=False)
    out = []
    for c in clouds:
        out += [c.xyz, c.fvec, c.xyz4]
    return out + [clouds[0].cam]


def _cloud_build_fake(fmaps, depths, intrs, extrs, stride, levels, depth_avg):
    V, S, C, H, W = fmaps.shape
    out = []
    for l in sorted(set(int(x) for x in levels)):
        P = V * (H >> l) * (W >> l)
        out += [fmaps.new_empty((S, P, 3), dtype=torch.float32), fmaps.new_empty((S, P, C)),
                fmaps.new_empty((S, P, 4), dtype=torch.float32)]
    return out + [fmaps.new_empty((V * S, ops._lib.CAM_STRIDE), dtype=torch.float32)]


# ---- knn: raster search over one level's records -------------------------------------------------------------------------------
def _knn_cuda(xyz4, cam, query, k, V, H, W, stride):
    c = ops.Cloud(0, None, None, None, None, xyz4, cam, V, H, W, float(stride))
    dist, idx = ops.knn(c, query, k, idx_dtype=torch.int64)
    return dist, idx


def _knn_fake(xyz4, cam, query, k, V, H, W, stride):
    B, M, _ = query.shape
    return query.new_empty((B, M, k), dtype=torch.float32), query.new_empty((B, M, k), dtype=torch.int64)


# ---- knn_corr: all levels fused; H, W, stride are level 0's (level i of the lists is i times halved) ---------------------------
def _knn_corr_cuda(xyz4, fvec, cam, targets, coords, k, V, H, W, stride):
    clouds = [ops.Cloud(i, None, f, None, None, x, cam, V, H >> i, W >> i, float(stride) * (1 << i)) for i, (x, f) in
              enumerate(zip(xyz4, fvec))]
    return ops.knn_corr([(c, c.fvec, k) for c in clouds], targets, coords, defer=False)


def _knn_corr_fake(xyz4, fvec, cam, targets, coords, k, V, H, W, stride):
    B, M, _ = targets.shape
    return targets.new_empty((B, M, len(xyz4) * k * 4), dtype=torch.float32)


# ---- track_input: corr [S, N, Lc] is copied into its columns (the in-place variant is ops.track_input(corr_in_out=True)) -------
def _track_input_cuda(coords, ffeats, track_mask, vis, pos_embed, time_embed, corr, flow_embed_dim):
    S, N, C = ffeats.shape
    Lc = corr.shape[-1]
    D = 3 * flow_embed_dim + 3 + Lc + C + 2
    out = torch.empty(N, S, D, device=ffeats.device, dtype=torch.float32)
    ops.corr_columns(out, flow_embed_dim, Lc).copy_(corr)
    return ops.track_input(coords, ffeats, track_mask, vis, pos_embed, time_embed, flow_embed_dim, Lc, out=out, corr_in_out=True)


def _track_input_fake(coords, ffeats, track_mask, vis, pos_embed, time_embed, corr, flow_embed_dim):
    S, N, C = ffeats.shape
    return ffeats.new_empty((N, S, 3 * flow_embed_dim + 3 + corr.shape[-1] + C + 2), dtype=torch.float32)


for _name, _cuda, _fake in (("cloud_build", _cloud_build_cuda, _cloud_build_fake), ("knn", _knn_cuda, _knn_fake),
                            ("knn_corr", _knn_corr_cuda, _knn_corr_fake), ("track_input", _track_input_cuda, _track_input_fake)):
    _lib_def.impl(_name, _cuda, "CUDA")
    torch.library.register_fake(f"mvt::{_name}", _fake, lib=_lib_def)


def cloud_build(fmaps, depths, intrs, extrs, stride=4.0, levels=(0, 1, 2, 3), depth_avg=False) -> List[torch.Tensor]:
    return torch.ops.mvt.cloud_build(fmaps, depths, intrs, extrs, float(stride), list(levels), bool(depth_avg))
