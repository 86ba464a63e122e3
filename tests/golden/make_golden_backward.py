"""Golden vectors for row f2 (backward of the path), produced by torch autograd through the REAL reference here:

    python tests/golden/make_golden_backward.py          # needs /root/reference (build container only)

  PointcloudCorrBlock.corr_sample (mvtracker/models/core/mvtracker/mvtracker.py:800-846): grads w.r.t. targets and fvec of
      loss = sum(out * w) for a seeded weight tensor w  -> bwdcorr_*.npz
  init_pointcloud_from_rgbd (mvtracker/models/core/model_utils.py:420-482): grad w.r.t. fmaps of sum(fvec * w) -> bwdcloud_*.npz
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ref_import  # noqa: E402
from mvtracker_b200 import synth  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def main():
    ref = ref_import.load()
    for name, B, N, M, C, k, G, add_xyz, seed in (("a", 2, 300, 37, 32, 8, 1, False, 21), ("b", 1, 500, 50, 128, 16, 1, False, 22),
                                                  ("c", 2, 200, 20, 32, 4, 4, True, 23)):
        g = torch.Generator().manual_seed(seed)
        xyz = torch.rand(B, N, 3, generator=g) * 2 - 1
        coords = xyz[:, torch.randperm(N, generator=g)[:M]] + 0.03 * torch.randn(B, M, 3, generator=g)
        fvec = torch.randn(B, N, C, generator=g, requires_grad=True)
        targets = torch.randn(B, M, C, generator=g, requires_grad=True)
        blk = ref.PointcloudCorrBlock(k=k, groups=G, xyz=xyz, fvec=fvec, corr_add_neighbor_offset=True, corr_add_neighbor_xyz=add_xyz)
        out = blk.corr_sample(targets, coords)
        w = torch.randn(out.shape, generator=g)
        (out * w).sum().backward()
        np.savez_compressed(os.path.join(OUT, f"bwdcorr_{name}.npz"), xyz=xyz.numpy(), fvec=fvec.detach().numpy(),
                            targets=targets.detach().numpy(), coords=coords.numpy(), w=w.numpy(), out=out.detach().numpy(),
                            grad_targets=targets.grad.numpy(), grad_fvec=fvec.grad.numpy(), k=np.int32(k), groups=np.int32(G),
                            add_xyz=np.bool_(add_xyz))
        print("bwdcorr", name, tuple(out.shape))
    for name, V, S, C, Hs, Ws, level, seed in (("a", 2, 2, 8, 13, 18, 0, 31), ("b", 2, 1, 4, 13, 22, 2, 32), ("c", 1, 2, 32, 8, 16, 3, 33)):
        sc = synth.make_scene(V, Hs * 4, Ws * 4, S=S, C=C, stride=4, zero_frac=0.1, seed=seed)
        fmaps = sc.fmaps.clone().requires_grad_(True)
        xyz, fvec = ref.init_pointcloud_from_rgbd(fmaps, sc.depths, sc.intrs, sc.extrs, stride=4, level=level)
        w = torch.randn(fvec.shape, generator=torch.Generator().manual_seed(seed + 1))
        (fvec * w).sum().backward()
        np.savez_compressed(os.path.join(OUT, f"bwdcloud_{name}.npz"), w=w.numpy(), grad_fmaps=fmaps.grad[0].numpy(),
                            V=np.int32(V), H=np.int32(Hs), W=np.int32(Ws), level=np.int32(level))
        print("bwdcloud", name, tuple(fmaps.grad.shape))


if __name__ == "__main__":
    main()
