"""Preprocessing of EvaluationPredictor.forward, the step immediately before the path (SURVEY.md 8 row f3;
mvtracker/models/evaluation_predictor_3dpt.py:72-120): nearest resize of rgb / depth to the model's input shape with the
matching intrinsics rescale, and the support grid -- a size x size grid of pixels per view at a few frames, lifted to world
space through the bilinearly sampled depth (mvt_lift_pixels) and appended to the query points."""
from typing import Optional, Tuple

import torch
import torch.nn.functional as F

from . import ops


def resize_inputs(rgbs, depths, intrs, interp_shape: Optional[Tuple[int, int]]):
    """evaluation_predictor_3dpt.py:72-88.  rgbs [B,V,T,3,H,W], depths [B,V,T,1,H,W], intrs [B,V,T,3,3]."""
    if interp_shape is None:
        return rgbs, depths, intrs
    B, V, T, _, Hr, Wr = rgbs.shape
    H, W = interp_shape
    rgbs = F.interpolate(rgbs.reshape(-1, 3, Hr, Wr), (H, W), mode="nearest").reshape(B, V, T, 3, H, W)
    depths = F.interpolate(depths.reshape(-1, 1, Hr, Wr), (H, W), mode="nearest").reshape(B, V, T, 1, H, W)
    scale = torch.tensor([[W / Wr, 0, 0], [0, H / Hr, 0], [0, 0, 1]], device=intrs.device, dtype=intrs.dtype)
    return rgbs, depths, torch.einsum("ij,BVTjk->BVTik", scale, intrs)


def grid_pixels(size: int, extent: Tuple[int, int], device) -> torch.Tensor:
    """get_points_on_a_grid (mvtracker/models/core/model_utils.py:361-417) with the default centre: [size*size, 2] (x, y)."""
    H, W = extent
    if size == 1:
        return torch.tensor([[W / 2, H / 2]], device=device)
    margin = W / 64
    gy, gx = torch.meshgrid(torch.linspace(margin, H - margin, size, device=device),
                            torch.linspace(margin, W - margin, size, device=device), indexing="ij")
    return torch.stack([gx, gy], dim=-1).reshape(-1, 2)


def support_grid_points(depths, intrs, extrs, grid_size: int, n_grids_per_view: int = 1) -> torch.Tensor:
    """evaluation_predictor_3dpt.py:99-118 for B = 1: depths [1,V,T,1,H,W], intrs [1,V,T,3,3], extrs [1,V,T,3,4]
    -> [1, n_t * V * grid_size^2, 4] support points (t, x, y, z), frames t = 0, step, 2*step, ... outer, views inner."""
    B, V, T, _, H, W = depths.shape
    assert B == 1
    if grid_size <= 0:
        return torch.zeros(1, 0, 4, device=depths.device)
    cam = ops.camera_prep(intrs[0], extrs[0]).reshape(V, T, -1)
    pix = grid_pixels(grid_size, (H, W), depths.device)
    pts = []
    for t in range(0, T, max(1, T // n_grids_per_view)):
        for v in range(V):
            pts.append(ops.lift_pixels(depths[0, v, t, 0], cam[v, t], pix, t=float(t)))
    return torch.cat(pts, dim=0)[None]
