"""GPU parity of the backward kernels: mvt_knn_corr_backward / mvt_cloud_build_backward against the
golden gradients (torch autograd through the real reference) and the oracle, and the autograd wiring of the drop-in."""
import numpy as np
import pytest
import torch

from mvtracker_b200 import dropin, ops, synth
from oracle import oracle
from parity import assert_close, golden

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _t(a, dtype=None):
    t = torch.from_numpy(np.ascontiguousarray(a)).to(DEV)
    return t if dtype is None else t.to(dtype)


@pytest.mark.parametrize("name,g", golden("bwdcorr"))
def test_corr_backward_matches_reference_golden(name, g):
    k, G = int(g["k"]), int(g["groups"])
    xyz, fvec, targets, coords = _t(g["xyz"]), _t(g["fvec"]), _t(g["targets"]), _t(g["coords"])
    grid = ops.build_grids([(xyz, None, None)], k_hint=k)[0]
    out, idx = ops.knn_corr([(grid, fvec, k)], targets, coords, groups=G, add_offset=True, add_xyz=bool(g["add_xyz"]), return_idx=True)
    D = g["w"].shape[-1]
    assert_close(out.reshape(g["out"].shape).cpu().numpy(), g["out"], 1e-4, name)
    gt, gf = ops.knn_corr_backward(fvec, targets, idx[0], _t(g["w"]), groups=G)
    assert_close(gt.cpu().numpy(), g["grad_targets"], 1e-4, f"{name} grad_targets")
    assert_close(gf.cpu().numpy(), g["grad_fvec"], 1e-4, f"{name} grad_fvec")
    # the gradient of a column slice of a wider (fused-levels) block: strided grad_out, accumulation into grad_targets
    wide = torch.zeros(*g["w"].shape[:2], 3 * k * D, device=DEV)
    wide[:, :, k * D:2 * k * D] = _t(g["w"]).reshape(*g["w"].shape[:2], k * D)
    view = wide[:, :, k * D:2 * k * D].unflatten(2, (k, D))
    gt2, gf2 = ops.knn_corr_backward(fvec, targets, idx[0], view, groups=G, grad_targets=gt.clone(), grad_fvec=gf.clone())
    assert_close(gt2.cpu().numpy(), 2 * g["grad_targets"], 1e-4, f"{name} accumulated grad_targets")
    assert_close(gf2.cpu().numpy(), 2 * g["grad_fvec"], 1e-4, f"{name} accumulated grad_fvec")


@pytest.mark.parametrize("name,g", golden("bwdcloud"))
def test_cloud_backward_matches_reference_golden(name, g):
    V, H, W, level = int(g["V"]), int(g["H"]), int(g["W"]), int(g["level"])
    got = ops.cloud_build_backward(_t(g["w"]), V, H, W, level)
    assert np.array_equal(got.cpu().numpy(), g["grad_fmaps"])          # a copy scaled by a power of two: bit-exact
    twice = ops.cloud_build_backward(_t(g["w"]), V, H, W, level, grad_fmaps=got.clone())
    assert np.array_equal(twice.cpu().numpy(), 2 * g["grad_fmaps"])


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_corr_backward_on_raster_cloud_vs_oracle(dtype):
    S, N, K = 3, 301, 16
    sc = synth.make_scene(3, 40 * 4, 56 * 4, S=S, C=128, zero_frac=0.1, seed=61, device=DEV, dtype=dtype)
    clouds = ops.cloud_build(sc.fmaps[0], sc.depths[0, :, :, 0], sc.intrs[0], sc.extrs[0], levels=range(4), want_valid=True)
    coords, targets = synth.make_tracks(clouds[0].xyz[0], N, S, valid=clouds[0].valid[0], dtype=dtype)
    out, idx = ops.knn_corr([(c, c.fvec, K) for c in clouds], targets, coords, return_idx=True)
    w = torch.randn(out.shape, generator=torch.Generator().manual_seed(5)).to(DEV)
    gt = None
    for l, c in enumerate(clouds):
        gl = w[:, :, l * K * 4:(l + 1) * K * 4].unflatten(2, (K, 4))
        gt, gf = ops.knn_corr_backward(c.fvec, targets, idx[l], gl, grad_targets=gt)
        _, gf_o = oracle.corr_backward(c.fvec.float().cpu().numpy(), targets.float().cpu().numpy(), idx[l].cpu().numpy(), gl.cpu().numpy())
        assert_close(gf.cpu().numpy(), gf_o, 1e-4, f"grad_fvec level {l}")
    gt_o = sum(oracle.corr_backward(c.fvec.float().cpu().numpy(), targets.float().cpu().numpy(), idx[l].cpu().numpy(),
                                    w[:, :, l * K * 4:(l + 1) * K * 4].unflatten(2, (K, 4)).cpu().numpy())[0].astype(np.float64)
               for l, c in enumerate(clouds))
    assert_close(gt.cpu().numpy(), gt_o, 1e-4, "grad_targets over 4 levels")


def test_dropin_autograd_end_to_end():
    """fmaps.requires_grad + targets.requires_grad through the drop-in seam (init_pointcloud_from_rgbd -> PointcloudCorrBlock
    .corr_sample): gradients equal the oracle's chain (corr_backward -> unproject_backward) on the same neighbour lists."""
    S, N, K, lvl = 2, 64, 8, 1
    sc = synth.make_scene(2, 24 * 4, 32 * 4, S=S, C=32, zero_frac=0.05, seed=67, device=DEV)
    fmaps = sc.fmaps.clone().requires_grad_(True)
    dropin.clear_caches()
    xyz, fvec = dropin.init_pointcloud_from_rgbd(fmaps, sc.depths, sc.intrs, sc.extrs, stride=4, level=lvl)
    assert fvec.requires_grad
    coords, targets = synth.make_tracks(xyz[0], N, S)
    targets = targets[:, :, :32].contiguous().requires_grad_(True)
    blk = dropin.PointcloudCorrBlock(k=K, groups=1, xyz=xyz, fvec=fvec, corr_add_neighbor_offset=True, corr_add_neighbor_xyz=False)
    out = blk.corr_sample(targets, coords)
    with torch.no_grad():
        out_ng = blk.corr_sample(targets.detach(), coords)
    assert torch.equal(out.detach(), out_ng)
    w = torch.randn(out.shape, generator=torch.Generator().manual_seed(9)).to(DEV)
    (out * w).sum().backward()
    _, idx, _ = oracle.corr_sample(K, xyz.cpu().numpy(), fvec.detach().cpu().numpy(), targets.detach().cpu().numpy(), coords.cpu().numpy(),
                                   return_idx=True)
    gt_o, gf_o = oracle.corr_backward(fvec.detach().cpu().numpy(), targets.detach().cpu().numpy(), idx, w.cpu().numpy())
    gm_o = oracle.unproject_backward(gf_o, 2, 24, 32, lvl)
    assert_close(targets.grad.cpu().numpy(), gt_o, 1e-4, "grad targets")
    assert_close(fmaps.grad[0].cpu().numpy(), gm_o, 1e-4, "grad fmaps")
    dropin.clear_caches()
