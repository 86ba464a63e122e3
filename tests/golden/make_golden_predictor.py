"""Golden vectors for row f3 (predictor preprocessing), produced by the REAL reference's functions here:

    python tests/golden/make_golden_predictor.py          # needs /root/reference (build container only)

Restates mvtracker/models/evaluation_predictor_3dpt.py:72-118 (inline code of EvaluationPredictor.forward) on the reference's
own get_points_on_a_grid / bilinear_sample2d / to_homogeneous / from_homogeneous.  Output: predictor_a.npz.
"""
import os
import sys

import numpy as np
import torch
import torch.nn.functional as F

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ref_import  # noqa: E402
from mvtracker_b200 import synth  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def main():
    ref_import.load()
    from mvtracker.models.core.model_utils import bilinear_sample2d, get_points_on_a_grid
    from mvtracker.utils.basic import from_homogeneous, to_homogeneous
    V, T, Hr, Wr, H, W, grid_size, n_grids = 2, 4, 30, 40, 18, 24, 5, 2
    sc = synth.make_scene(V, Hr, Wr, S=T, C=3, stride=1, zero_frac=0.05, seed=41)
    rgbs, depths, intrs, extrs = sc.fmaps, sc.depths, sc.intrs, sc.extrs
    B = 1
    raw = dict(rgbs=rgbs[0].numpy(), depths=depths[0].numpy(), intrs=intrs[0].numpy(), extrs=extrs[0].numpy())
    # evaluation_predictor_3dpt.py:72-88
    rgbs = F.interpolate(rgbs.reshape(-1, 3, Hr, Wr), (H, W), mode="nearest").reshape(B, V, T, 3, H, W)
    depths = F.interpolate(depths.reshape(-1, 1, Hr, Wr), (H, W), mode="nearest").reshape(B, V, T, 1, H, W)
    tr = torch.tensor([[W / Wr, 0, 0], [0, H / Hr, 0], [0, 0, 1]], dtype=intrs.dtype)
    intrs = torch.einsum("ij,BVTjk->BVTik", tr, intrs)
    # :94-97
    intrs_inv = torch.inverse(intrs.float())
    extrs_square = torch.eye(4)[None].repeat(B, V, T, 1, 1)
    extrs_square[:, :, :, :3, :] = extrs
    extrs_inv = torch.inverse(extrs_square.float())
    # :101-117
    grid_points = []
    pixel_xy = get_points_on_a_grid(grid_size, (H, W))
    pixel_xy_homo = to_homogeneous(pixel_xy)
    for t in range(0, T, max(1, T // n_grids)):
        for view_idx in range(V):
            camera_z = bilinear_sample2d(depths[0, view_idx, t][None], pixel_xy[..., 0], pixel_xy[..., 1]).permute(0, 2, 1)
            camera_xyz = torch.einsum("Bij,BNj->BNi", intrs_inv[:, view_idx, t, :, :], pixel_xy_homo)
            camera_xyz = camera_xyz * camera_z
            world_xyz_homo = torch.einsum("Bij,BNj->BNi", extrs_inv[:, view_idx, t, :, :], to_homogeneous(camera_xyz))
            world_xyz = from_homogeneous(world_xyz_homo)
            grid_points.append(torch.cat([torch.ones_like(world_xyz[:, :, :1]) * t, world_xyz], dim=2))
    grid_points = torch.cat(grid_points, dim=1)
    np.savez_compressed(os.path.join(OUT, "predictor_a.npz"), **raw, interp_shape=np.array([H, W], np.int32),
                        grid_size=np.int32(grid_size), n_grids_per_view=np.int32(n_grids), rgbs_resized=rgbs[0].numpy(),
                        depths_resized=depths[0].numpy(), intrs_resized=intrs[0].numpy(), pixel_xy=pixel_xy[0].numpy(),
                        support_points=grid_points[0].numpy())
    print("predictor a", tuple(grid_points.shape))


if __name__ == "__main__":
    main()
