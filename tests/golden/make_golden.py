"""Generate the committed golden vectors by running the REAL reference (CPU, torch) here.

    python tests/golden/make_golden.py          # needs /root/reference (build container only)

Outputs (small .npz, committed): unproject_*.npz, knn_*.npz, corr_*.npz.  Each file stores
the seeded inputs and the outputs of the reference's own functions:
  init_pointcloud_from_rgbd  (mvtracker/models/core/model_utils.py:420-482)
  knn == _knn_torch          (mvtracker/models/core/mvtracker/mvtracker.py:75-79; pointops is absent)
  PointcloudCorrBlock.corr_sample (mvtracker/models/core/mvtracker/mvtracker.py:769-846)
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


def scene(V, S, C, Hs, Ws, stride, zero_frac, seed):
    sc = synth.make_scene(V, Hs * stride, Ws * stride, S=S, C=C, stride=stride, zero_frac=zero_frac, seed=seed)
    # perturb cameras over time so per-(v,s) matrices differ
    g = torch.Generator().manual_seed(seed + 1)
    sc.extrs[..., 3] += 0.01 * torch.randn(sc.extrs[..., 3].shape, generator=g)
    return sc


def gen_unproject(ref):
    cases = [
        dict(name="a", V=2, S=2, C=8, Hs=22, Ws=37, stride=4, zero=0.1, levels=[0, 1, 2, 3], mode="nearest"),
        dict(name="b", V=3, S=1, C=3, Hs=16, Ws=24, stride=1, zero=0.0, levels=[0], mode="nearest"),
        dict(name="c", V=2, S=2, C=4, Hs=17, Ws=18, stride=4, zero=0.3, levels=[1, 2], mode="avg"),
        dict(name="d", V=1, S=3, C=32, Hs=8, Ws=40, stride=4, zero=0.05, levels=[0, 3], mode="nearest"),
    ]
    for cs in cases:
        sc = scene(cs["V"], cs["S"], cs["C"], cs["Hs"], cs["Ws"], cs["stride"], cs["zero"], seed=100 + ord(cs["name"]))
        d = dict(fmaps=sc.fmaps[0].numpy(), depths=sc.depths[0, :, :, 0].numpy(), intrs=sc.intrs[0].numpy(),
                 extrs=sc.extrs[0].numpy(), stride=np.int32(cs["stride"]), levels=np.array(cs["levels"], np.int32),
                 mode=np.array(cs["mode"]))
        for lvl in cs["levels"]:
            xyz, fvec, valid = ref.init_pointcloud_from_rgbd(sc.fmaps, sc.depths, sc.intrs, sc.extrs, stride=cs["stride"],
                                                             level=lvl, depth_interp_mode=cs["mode"],
                                                             return_validity_mask=True)
            d[f"xyz_l{lvl}"] = xyz.numpy()
            d[f"fvec_l{lvl}"] = fvec.numpy()
            d[f"valid_l{lvl}"] = valid.numpy()
        np.savez_compressed(os.path.join(OUT, f"unproject_{cs['name']}.npz"), **d)
        print("unproject", cs["name"], {k: v.shape for k, v in d.items() if k.startswith("xyz")})


def cloud_and_queries(seed, N, M, B, offset=0.0, dup=0):
    g = torch.Generator().manual_seed(seed)
    # a bumpy surface sheet plus volume noise, like a fused depth cloud
    uv = torch.rand(B, N, 2, generator=g) * 4 - 2
    z = 0.3 * torch.sin(2 * uv[..., :1]) * torch.cos(3 * uv[..., 1:]) + 0.01 * torch.randn(B, N, 1, generator=g)
    ref_xyz = torch.cat([uv, z], -1) + offset
    if dup:
        ref_xyz[:, 10:10 + dup] = ref_xyz[:, 10:11]          # identical points (depth-0 pixels collapse)
    sel = torch.randint(0, N, (B, M), generator=g)
    q = torch.gather(ref_xyz, 1, sel[..., None].expand(B, M, 3)) + 0.03 * torch.randn(B, M, 3, generator=g)
    q[:, :2] = ref_xyz[:, 10:12]                               # two queries exactly on cloud points (the dup cluster)
    q[:, 2] = q[:, 2] + 5.0                                    # one query far outside the cloud
    return ref_xyz.contiguous(), q.contiguous()


def gen_knn(ref):
    cases = [
        dict(name="a", N=700, M=48, B=2, k=16, offset=0.0, dup=0),
        dict(name="b", N=700, M=48, B=2, k=8, offset=0.0, dup=40),
        dict(name="c", N=300, M=32, B=3, k=1, offset=0.0, dup=5),
        dict(name="d", N=900, M=40, B=1, k=16, offset=10.0, dup=0),   # far from origin: cdist expansion error
    ]
    for cs in cases:
        xyz, q = cloud_and_queries(200 + ord(cs["name"]), cs["N"], cs["M"], cs["B"], cs["offset"], cs["dup"])
        dists, idx = ref.knn(cs["k"], xyz, q)
        np.savez_compressed(os.path.join(OUT, f"knn_{cs['name']}.npz"), xyz=xyz.numpy(), query=q.numpy(),
                            k=np.int32(cs["k"]), dists=dists.numpy(), idx=idx.numpy().astype(np.int32))
        print("knn", cs["name"], tuple(idx.shape))


def gen_corr(ref):
    cases = [
        dict(name="a", N=600, M=40, B=2, C=32, k=16, G=1, off=True, xyz=False),
        dict(name="b", N=600, M=40, B=2, C=32, k=8, G=4, off=True, xyz=True),
        dict(name="c", N=500, M=24, B=1, C=128, k=16, G=1, off=False, xyz=False),
        dict(name="d", N=500, M=24, B=3, C=128, k=16, G=2, off=True, xyz=False),
    ]
    for cs in cases:
        xyz, q = cloud_and_queries(300 + ord(cs["name"]), cs["N"], cs["M"], cs["B"])
        g = torch.Generator().manual_seed(400 + ord(cs["name"]))
        fvec = torch.randn(cs["B"], cs["N"], cs["C"], generator=g)
        targets = torch.randn(cs["B"], cs["M"], cs["C"], generator=g)
        blk = ref.PointcloudCorrBlock(k=cs["k"], groups=cs["G"], xyz=xyz, fvec=fvec, corr_add_neighbor_offset=cs["off"],
                                      corr_add_neighbor_xyz=cs["xyz"])
        out = blk.corr_sample(targets, q)
        np.savez_compressed(os.path.join(OUT, f"corr_{cs['name']}.npz"), xyz=xyz.numpy(), fvec=fvec.numpy(),
                            targets=targets.numpy(), coords=q.numpy(), k=np.int32(cs["k"]), groups=np.int32(cs["G"]),
                            add_offset=np.bool_(cs["off"]), add_xyz=np.bool_(cs["xyz"]), out=out.numpy())
        print("corr", cs["name"], tuple(out.shape))


if __name__ == "__main__":
    torch.manual_seed(72)
    np.random.seed(72)
    ref = ref_import.load()
    gen_unproject(ref)
    gen_knn(ref)
    gen_corr(ref)
