"""GPU parity: the CUDA path (through the C ABI) against the CPU oracle and the reference-generated golden vectors.

Bars (north_star): neighbour indices bit-exact against the canonical oracle (ties by index); unprojected points and
correlation volumes within 1e-4 relative (fp32), 2e-2 with bf16 features."""
import numpy as np
import pytest
import torch

from mvtracker_b200 import dropin, ops, synth
from oracle import oracle
from parity import assert_close, golden, knn_sets_equivalent

pytestmark = pytest.mark.gpu
DEV = "cuda"


def t(a, dtype=torch.float32):
    return torch.as_tensor(np.ascontiguousarray(a)).to(DEV, dtype)


# ---------------------------------------------------------------------------------------------------------------
# cloud build
# ---------------------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name,g", golden("unproject"))
def test_cloud_build_matches_golden(name, g):
    levels = g["levels"].tolist()
    clouds = ops.cloud_build(t(g["fmaps"]), t(g["depths"]), t(g["intrs"]), t(g["extrs"]), stride=int(g["stride"]),
                             levels=levels, depth_interp_mode=str(g["mode"]), want_valid=True)
    for c in clouds:
        assert_close(c.xyz.cpu().numpy(), g[f"xyz_l{c.level}"], 1e-4, f"{name} xyz l{c.level}")
        assert_close(c.fvec.cpu().numpy(), g[f"fvec_l{c.level}"], 1e-5, f"{name} fvec l{c.level}")
        assert (c.valid.cpu().numpy() == g[f"valid_l{c.level}"]).all()


@pytest.mark.parametrize("V,S,C,Hs,Ws,stride", [(3, 2, 128, 45, 80, 4), (2, 3, 32, 33, 70, 4), (1, 1, 3, 19, 21, 1),
                                                (2, 2, 6, 8, 32, 4)])
def test_cloud_build_matches_oracle(V, S, C, Hs, Ws, stride):
    sc = synth.make_scene(V, Hs * stride, Ws * stride, S=S, C=C, stride=stride, zero_frac=0.2, seed=5, device=DEV)
    fm, dp, K, E = sc.fmaps[0], sc.depths[0, :, :, 0], sc.intrs[0], sc.extrs[0]
    levels = [l for l in range(4) if (Hs >> l) > 0 and (Ws >> l) > 0]
    clouds = ops.cloud_build(fm, dp, K, E, stride=stride, levels=levels, want_valid=True)
    for c in clouds:
        xyz, fvec, valid = oracle.unproject(fm.cpu().numpy(), dp.cpu().numpy(), K.cpu().numpy(), E.cpu().numpy(),
                                            stride=stride, level=c.level, return_validity_mask=True)
        assert_close(c.xyz.cpu().numpy(), xyz, 1e-4, f"xyz l{c.level}")
        assert_close(c.fvec.cpu().numpy(), fvec, 1e-5, f"fvec l{c.level}")
        assert (c.valid.cpu().numpy() == valid).all()
        # spacing hint = mean of depth*stride_l/fx over valid pixels
        sp = c.spacing.cpu().numpy()
        assert (sp[:, 1] == valid.sum(1)).all()
    # one level at a time gives the same tensors as the fused pass
    single = ops.cloud_build(fm, dp, K, E, stride=stride, levels=[levels[-1]])[0]
    assert torch.equal(single.xyz, clouds[-1].xyz) and torch.equal(single.fvec, clouds[-1].fvec)


def test_cloud_build_bf16_features():
    sc = synth.make_scene(2, 64 * 4, 96 * 4, S=2, C=128, zero_frac=0.1, seed=6, device=DEV)
    fm = sc.fmaps[0].to(torch.bfloat16)
    clouds = ops.cloud_build(fm, sc.depths[0, :, :, 0], sc.intrs[0], sc.extrs[0], levels=[0, 1, 2, 3])
    ref = ops.cloud_build(fm.float(), sc.depths[0, :, :, 0], sc.intrs[0], sc.extrs[0], levels=[0, 1, 2, 3])
    for c, r in zip(clouds, ref):
        assert c.fvec.dtype == torch.bfloat16
        assert torch.equal(c.xyz, r.xyz)
        err = (c.fvec.float() - r.fvec).abs().max().item()
        assert err <= 2e-2 * r.fvec.abs().max().item(), f"bf16 fvec l{c.level}: {err}"
    assert torch.equal(clouds[0].fvec, ref[0].fvec.to(torch.bfloat16))   # level 0 is a pure transpose


# ---------------------------------------------------------------------------------------------------------------
# kNN
# ---------------------------------------------------------------------------------------------------------------
def check_knn_exact(xyz, q, k, mask=None, spacing=None, cell_scale=0.0, k_hint=None):
    grid = ops.build_grids([(xyz, mask, spacing)], k_hint=k_hint or k, cell_scale=cell_scale)[0]
    dist, idx = ops.knn(grid, q, k, idx_dtype=torch.int32)
    od, oi = oracle.knn(k, xyz.cpu().numpy(), q.cpu().numpy(), None if mask is None else mask.cpu().numpy())
    idx, dist = idx.cpu().numpy(), dist.cpu().numpy()
    assert (idx == oi).all(), f"{(idx != oi).any(-1).sum()} / {idx.shape[0] * idx.shape[1]} queries differ"
    assert (dist == od).all()
    return grid


def check_knn_raster(cloud, q, k, skip_invalid=False):
    """Raster search (no index) against the canonical oracle on the same CUDA-built cloud: bit exact."""
    dist, idx = ops.knn(ops.Search.from_cloud(cloud, skip_invalid=skip_invalid, tiles=True), q, k, idx_dtype=torch.int32)
    mask = cloud.valid.cpu().numpy() if skip_invalid else None
    od, oi = oracle.knn(k, cloud.xyz.cpu().numpy(), q.cpu().numpy(), mask)
    idx, dist = idx.cpu().numpy(), dist.cpu().numpy()
    bad = (idx != oi).any(-1)
    assert not bad.any(), f"{bad.sum()} / {bad.size} queries differ, first at {np.argwhere(bad)[0]}"
    assert (dist == od).all()


@pytest.mark.parametrize("name,g", golden("knn"))
def test_knn_golden(name, g):
    k = int(g["k"])
    xyz, q = t(g["xyz"]), t(g["query"])
    check_knn_exact(xyz, q, k)
    # and, as sets, against the reference's own output (ties / cdist error adjudicated in f64)
    grid = ops.build_grids([(xyz, None, None)], k_hint=k)[0]
    _, idx = ops.knn(grid, q, k)
    assert idx.dtype == torch.int64
    knn_sets_equivalent(idx.cpu().numpy(), g["idx"], g["xyz"], g["query"], name)


@pytest.mark.parametrize("k", [1, 8, 16, 32])
@pytest.mark.parametrize("cell_scale", [0.0, 0.3, 4.0])
def test_knn_scene_exact(k, cell_scale):
    """Fused multi-view cloud incl. depth-0 pixels (thousands of identical points at each camera centre)."""
    sc = synth.make_scene(3, 45 * 4, 80 * 4, S=2, C=4, zero_frac=0.3, seed=11, device=DEV)
    c = ops.cloud_build(sc.fmaps[0], sc.depths[0, :, :, 0], sc.intrs[0], sc.extrs[0], levels=[0], want_valid=True)[0]
    coords, _ = synth.make_tracks(c.xyz[0], 300, 2, C=4, valid=c.valid[0])
    coords[:, :4] += 3.0                         # far from every surface -> ring fallback
    coords[:, 4] = c.xyz[:, 0]                    # exactly on a cloud point
    cam_centre = c.xyz[0][~c.valid[0]][0]
    coords[:, 5] = cam_centre                     # on the depth-0 pile: all-ties, must resolve by index
    coords[:, 6] = cam_centre + 1e-3
    check_knn_exact(c.xyz, coords, k, spacing=c.spacing, cell_scale=cell_scale)
    check_knn_exact(c.xyz, coords, k, spacing=None, cell_scale=cell_scale)      # measure-based cell size
    if cell_scale == 0.0:
        check_knn_raster(c, coords, k)


def test_knn_mask_and_degenerate_clouds():
    g = torch.Generator(device=DEV).manual_seed(3)
    xyz = torch.rand(2, 5000, 3, device=DEV, generator=g)
    q = torch.rand(2, 200, 3, device=DEV, generator=g)
    mask = torch.rand(2, 5000, device=DEV, generator=g) > 0.5
    check_knn_exact(xyz, q, 16, mask=mask)
    # planar cloud (one collapsed axis), all-identical cloud, tiny cloud (exhaustive path), k == N
    flat = xyz.clone(); flat[..., 2] = 0.25
    check_knn_exact(flat, q, 8)
    same = torch.ones(1, 2000, 3, device=DEV)
    check_knn_exact(same, q[:1], 16)
    check_knn_exact(xyz[:, :40].contiguous(), q, 16)
    check_knn_exact(xyz[:, :16].contiguous(), q, 16)
    # far-from-origin scene (the case where the reference's cdist loses neighbours)
    check_knn_exact(xyz + 100.0, q + 100.0, 16)


def test_knn_fewer_points_than_k():
    xyz = torch.rand(1, 2000, 3, device=DEV)
    mask = torch.zeros(1, 2000, dtype=torch.bool, device=DEV); mask[0, :5] = True
    grid = ops.build_grids([(xyz, mask, None)], k_hint=8)[0]
    dist, idx = ops.knn(grid, torch.rand(1, 7, 3, device=DEV), 8)
    assert (idx[..., 5:] == -1).all() and torch.isinf(dist[..., 5:]).all() and (idx[..., :5] >= 0).all()


# ---------------------------------------------------------------------------------------------------------------
# fused kNN + gather + correlation
# ---------------------------------------------------------------------------------------------------------------
def check_corr(xyz, fvec, targets, coords, k, G=1, add_offset=True, add_xyz=False, tol=1e-4, spacing=None, cloud=None):
    where = cloud if cloud is not None else ops.build_grids([(xyz, None, spacing)], k_hint=k)[0]
    out, idx = ops.knn_corr([(where, fvec, k)], targets, coords, groups=G, add_offset=add_offset, add_xyz=add_xyz,
                            return_idx=True)
    D = G + 3 * add_offset + 3 * add_xyz
    out = out.reshape(xyz.shape[0], coords.shape[1], k, D).cpu().numpy()
    o_out, o_idx, _ = oracle.corr_sample(k, xyz.cpu().numpy(), fvec.float().cpu().numpy(), targets.float().cpu().numpy(),
                                         coords.cpu().numpy(), G, add_offset, add_xyz, return_idx=True)
    assert (idx[0].cpu().numpy() == o_idx).all()
    assert_close(out[..., :G], o_out[..., :G], tol, "corr")
    assert (out[..., G:] == o_out[..., G:]).all(), "offsets / xyz must be bit exact"
    return out


@pytest.mark.parametrize("name,g", golden("corr"))
def test_corr_golden(name, g):
    k, G = int(g["k"]), int(g["groups"])
    out = check_corr(t(g["xyz"]), t(g["fvec"]), t(g["targets"]), t(g["coords"]), k, G, bool(g["add_offset"]),
                     bool(g["add_xyz"]))
    ref = g["out"]
    same = np.isclose(out, ref, rtol=1e-4, atol=1e-5).reshape(-1, k * ref.shape[-1]).all(-1)
    assert same.mean() >= 0.95, f"{name}: only {same.mean():.3f} of queries match the reference positionally"


@pytest.mark.parametrize("C,G,k,add_xyz", [(128, 1, 16, False), (128, 1, 8, True), (128, 4, 16, False), (32, 1, 16, False),
                                           (32, 4, 8, True), (6, 3, 4, False), (128, 1, 32, False)])
def test_corr_scene(C, G, k, add_xyz):
    sc = synth.make_scene(2, 40 * 4, 64 * 4, S=3, C=C, zero_frac=0.1, seed=21, device=DEV)
    c = ops.cloud_build(sc.fmaps[0], sc.depths[0, :, :, 0], sc.intrs[0], sc.extrs[0], levels=[0], want_valid=True)[0]
    coords, targets = synth.make_tracks(c.xyz[0], 257, 3, C=C, valid=c.valid[0])
    check_corr(c.xyz, c.fvec, targets, coords, k, G, True, add_xyz, spacing=c.spacing)
    check_corr(c.xyz, c.fvec, targets, coords, k, G, True, add_xyz, cloud=c)


def test_corr_bf16():
    sc = synth.make_scene(2, 40 * 4, 64 * 4, S=2, C=128, zero_frac=0.1, seed=22, device=DEV)
    c = ops.cloud_build(sc.fmaps[0].to(torch.bfloat16), sc.depths[0, :, :, 0], sc.intrs[0], sc.extrs[0], levels=[0],
                        want_valid=True)[0]
    coords, targets = synth.make_tracks(c.xyz[0], 200, 2, C=128, valid=c.valid[0], dtype=torch.bfloat16)
    check_corr(c.xyz, c.fvec, targets, coords, 16, tol=2e-2, spacing=c.spacing)
    check_corr(c.xyz, c.fvec, targets, coords, 16, tol=2e-2, cloud=c)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("c0", [195, 192])     # 195: the reference's column (after the 195 flow-embedding floats); 192: aligned
def test_corr_written_into_transformer_input_layout(dtype, c0):
    """Row f1: the [S, N, 256] block written straight into a [N, S, 581] transformer-input buffer at its column offset
    (mvtracker.py:374-384 builds that with permute + cat): frame / track strides of the output are free."""
    S, N, K = 3, 157, 16
    sc = synth.make_scene(3, 40 * 4, 56 * 4, S=S, C=128, zero_frac=0.1, seed=81, device=DEV, dtype=dtype)
    clouds = ops.cloud_build(sc.fmaps[0], sc.depths[0, :, :, 0], sc.intrs[0], sc.extrs[0], levels=range(4), want_valid=True)
    coords, targets = synth.make_tracks(clouds[0].xyz[0], N, S, valid=clouds[0].valid[0], dtype=dtype)
    lv = [(c, c.fvec, K) for c in clouds]
    dense = ops.knn_corr(lv, targets, coords)
    tin = torch.full((N, S, 581), -7.0, device=DEV)
    view = tin.permute(1, 0, 2)[:, :, c0:c0 + 256]
    got = ops.knn_corr(lv, targets, coords, out=view)
    assert got.data_ptr() == view.data_ptr()
    assert torch.equal(tin[:, :, c0:c0 + 256], dense.permute(1, 0, 2))
    assert (tin[:, :, :c0] == -7.0).all() and (tin[:, :, c0 + 256:] == -7.0).all()      # nothing else is touched


def test_corr_fused_levels_equal_separate_launches():
    sc = synth.make_scene(3, 48 * 4, 64 * 4, S=4, C=128, zero_frac=0.05, seed=23, device=DEV)
    clouds = ops.cloud_build(sc.fmaps[0], sc.depths[0, :, :, 0], sc.intrs[0], sc.extrs[0], levels=[0, 1, 2, 3], want_valid=True)
    grids = ops.build_grids([(c.xyz, None, c.spacing) for c in clouds], k_hint=16)
    coords, targets = synth.make_tracks(clouds[0].xyz[0], 333, 4, valid=clouds[0].valid[0])
    fused = ops.knn_corr([(g, c.fvec, 16) for g, c in zip(grids, clouds)], targets, coords)
    assert fused.shape == (4, 333, 4 * 16 * 4)
    fused_r = ops.knn_corr([(c, c.fvec, 16) for c in clouds], targets, coords)
    assert torch.equal(fused, fused_r), "raster and grid search must give identical outputs"
    for i, (g, c) in enumerate(zip(grids, clouds)):
        one = ops.knn_corr([(g, c.fvec, 16)], targets, coords)
        assert torch.equal(one, fused[:, :, i * 64:(i + 1) * 64])
        o_out = oracle.corr_sample(16, c.xyz.cpu().numpy(), c.fvec.cpu().numpy(), targets.cpu().numpy(), coords.cpu().numpy())
        assert_close(one.reshape(4, 333, 16, 4).cpu().numpy(), o_out, 1e-4, f"level {i}")


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("k", [16, 8])
def test_corr_half_warp_groups_identical(dtype, k):
    """mvt_debug_set_group_width(16): two queries per warp in lock step; outputs must equal the warp-per-query kernel
    bit for bit (odd M -> the last half-warp group of every frame is padding)."""
    from mvtracker_b200 import _lib
    sc = synth.make_scene(3, 48 * 4, 64 * 4, S=3, C=128, zero_frac=0.2, seed=29, device=DEV, dtype=dtype)
    clouds = ops.cloud_build(sc.fmaps[0], sc.depths[0, :, :, 0], sc.intrs[0], sc.extrs[0], levels=[0, 1, 2, 3], want_valid=True)
    coords, targets = synth.make_tracks(clouds[0].xyz[0], 301, 3, valid=clouds[0].valid[0], dtype=dtype, jitter=0.05)
    coords[:, :7] += 3.0          # a few queries far from every surface (many passes)
    lv = [(c, c.fvec, k) for c in clouds]
    L = _lib.lib()
    try:
        _lib.check(L.mvt_debug_set_group_width(32), "set_group_width")
        a, ia = ops.knn_corr(lv, targets, coords, return_idx=True)
        _lib.check(L.mvt_debug_set_group_width(16), "set_group_width")
        b, ib = ops.knn_corr(lv, targets, coords, return_idx=True)
    finally:
        L.mvt_debug_set_group_width(32)
    assert torch.equal(ia, ib)
    assert torch.equal(a, b)
    _, oi = oracle.knn(k, clouds[0].xyz.cpu().numpy(), coords.cpu().numpy())
    assert (ib[0].cpu().numpy() == oi).all()


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_corr_warm_start_is_exact(dtype):
    """idx_hint (previous iteration's neighbours, or garbage) only bounds the first radius: outputs must not change."""
    sc = synth.make_scene(3, 48 * 4, 64 * 4, S=3, C=128, zero_frac=0.2, seed=37, device=DEV, dtype=dtype)
    clouds = ops.cloud_build(sc.fmaps[0], sc.depths[0, :, :, 0], sc.intrs[0], sc.extrs[0], levels=[0, 1, 2, 3], want_valid=True)
    coords, targets = synth.make_tracks(clouds[0].xyz[0], 257, 3, valid=clouds[0].valid[0], dtype=dtype, jitter=0.03)
    coords[:, :6] += 1.5
    lv = [(c, c.fvec, 16) for c in clouds]
    cold, idx0 = ops.knn_corr(lv, targets, coords, return_idx=True)
    g = torch.Generator(device=DEV).manual_seed(38)
    moved = coords + 0.02 * torch.randn(coords.shape, device=DEV, generator=g)
    want, widx = ops.knn_corr(lv, targets, moved, return_idx=True)
    # (a) hint = last iteration's neighbours, written back in place
    buf = idx0.clone()
    got = ops.knn_corr(lv, targets, moved, idx_hint=buf, idx_out=buf)
    assert torch.equal(got, want) and torch.equal(buf, widx)
    # (b) a useless hint: random points, some missing (-1), some out of range
    junk = torch.randint(-5, 400, idx0.shape, device=DEV, generator=g, dtype=torch.int32)
    junk[0, :, ::3] = 10 ** 9
    got, gidx = ops.knn_corr(lv, targets, moved, idx_hint=junk, return_idx=True)
    assert torch.equal(got, want) and torch.equal(gidx, widx)
    # (c) the window driver warm-starts by itself
    from mvtracker_b200.hotpath import WindowClouds
    win = WindowClouds(clouds, None, 16, 1, True, False)
    win.warm_start = True
    assert torch.equal(win.corr(targets, coords), cold)
    assert torch.equal(win.corr(targets, moved), want)
    win.warm_start = False
    assert torch.equal(win.corr(targets, moved), want)


def test_knn_gather_feature_init():
    sc = synth.make_scene(3, 40 * 4, 64 * 4, S=2, C=128, zero_frac=0.1, seed=24, device=DEV)
    c = ops.cloud_build(sc.fmaps[0], sc.depths[0, :, :, 0], sc.intrs[0], sc.extrs[0], levels=[0], want_valid=True)[0]
    coords, _ = synth.make_tracks(c.xyz[0], 100, 2, valid=c.valid[0])
    grid = ops.build_grids([(c.xyz, None, c.spacing)], k_hint=1)[0]
    _, oi = oracle.knn(1, c.xyz.cpu().numpy(), coords.cpu().numpy())
    for where in (grid, c):
        feat, idx = ops.knn_gather(where, c.fvec, coords, return_idx=True)
        assert (idx.cpu().numpy() == oi[..., 0]).all()
        expect = torch.gather(c.fvec, 1, idx.long()[..., None].expand(-1, -1, 128))
        assert torch.equal(feat, expect)
    # one frame of the window (what the feature-init loop asks for)
    feat1 = ops.knn_gather(ops.Search.from_cloud(c, frame=1, tiles=True), c.fvec[1:2], coords[1:2])
    assert torch.equal(feat1[0], expect[1])


# ---------------------------------------------------------------------------------------------------------------
# raster search: geometry edge cases (all must stay bit exact against the brute-force oracle)
# ---------------------------------------------------------------------------------------------------------------
def _edge_queries(c, sc, n=160):
    """Queries on and off the surfaces, outside every frustum, behind cameras, on camera centres and principal planes."""
    S = c.xyz.shape[0]
    coords, _ = synth.make_tracks(c.xyz[0], n, S, C=4, valid=c.valid[0], jitter=0.03)
    g = torch.Generator(device=DEV).manual_seed(5)
    E = sc.extrs[0, :, 0]                                     # [V,3,4]
    centres = -(E[:, :, :3].transpose(1, 2) @ E[:, :, 3:]).squeeze(-1)     # camera centres in world coordinates
    coords[:, 0:8] += 0.5 * torch.randn(S, 8, 3, device=DEV, generator=g)   # floating off the surfaces (big balls)
    coords[:, 8:12] *= 3.0                                    # far outside the scene
    coords[:, 12] = centres[0]                                # exactly a camera centre (the depth-0 pile)
    coords[:, 13] = centres[0] + 0.002
    coords[:, 14] = centres[-1] * 1.3                         # behind a camera
    coords[:, 15] = centres[0] + 0.3 * E[0, 0, :3]            # on camera 0's principal plane (z_cam == 0)
    coords[:, 16] = 0.0                                       # scene centre, inside the ground plane
    coords[:, 17] = torch.tensor([0.0, 0.0, 50.0], device=DEV)
    coords[:, 18] = c.xyz[:, 0]                               # exactly a cloud point
    return coords


@pytest.mark.parametrize("V,Hs,Ws,zero", [(3, 45, 80, 0.3), (1, 40, 64, 0.0), (8, 24, 40, 0.05), (2, 17, 23, 0.6)])
@pytest.mark.parametrize("k", [1, 16])
def test_raster_knn_edge_cases(V, Hs, Ws, zero, k):
    sc = synth.make_scene(V, Hs * 4, Ws * 4, S=2, C=4, zero_frac=zero, seed=40 + V, device=DEV)
    levels = [l for l in range(4) if (Hs >> l) * (Ws >> l) * V >= 16]
    clouds = ops.cloud_build(sc.fmaps[0], sc.depths[0, :, :, 0], sc.intrs[0], sc.extrs[0], levels=levels, want_valid=True)
    coords = _edge_queries(clouds[0], sc)
    for c in clouds:
        check_knn_raster(c, coords, k)
        check_knn_raster(c, coords, k, skip_invalid=True)     # "intended" filter_invalid: depth <= 0 not searchable


@pytest.mark.parametrize("V", [2, 4])
def test_raster_knn_queries_beside_and_behind_cameras(V):
    """Random queries all over the camera ring, searched in coarse rasters: the k-th distance is metres, balls reach
    behind / beside the cameras, and their arc of directions wraps past +-90 and 180 degrees (regression: an arc that
    started in front of a camera and ended past 180 degrees was declared invisible)."""
    sc = synth.make_scene(V, 48 * 4, 64 * 4, S=2, C=4, zero_frac=0.1, seed=60 + V, device=DEV)
    clouds = ops.cloud_build(sc.fmaps[0], sc.depths[0, :, :, 0], sc.intrs[0], sc.extrs[0], levels=[1, 2, 3], want_valid=True)
    g = torch.Generator(device=DEV).manual_seed(61)
    coords = (torch.rand(2, 1500, 3, device=DEV, generator=g) - 0.5) * torch.tensor([12.0, 12.0, 4.0], device=DEV)
    E = sc.extrs[0, :, 0]
    centres = -(E[:, :, :3].transpose(1, 2) @ E[:, :, 3:]).squeeze(-1)
    coords[:, :500] = centres[torch.arange(500, device=DEV) % V] + 0.8 * torch.randn(2, 500, 3, device=DEV, generator=g)
    for c in clouds:
        check_knn_raster(c, coords, 16)
        check_knn_raster(c, coords, 3, skip_invalid=True)


def test_tile_depth_tables_and_culled_search():
    """The per-tile depth ranges (mvt_tile_depth_build) against numpy, and the searches that use them -- queries next to
    the cameras and far off the surfaces, whose balls project onto large pixel boxes -- against the oracle and against
    the same search without tables; the statistics prove the tile path ran and culled."""
    import ctypes
    from mvtracker_b200 import _lib
    V, S = 5, 2
    sc = synth.make_scene(V, 80 * 4, 96 * 4, S=S, C=4, zero_frac=0.1, seed=90, device=DEV)
    sc.depths[0, 1, 0, 0, :64, :64] = float("nan")          # a block of NaN depths: never a neighbour, never in a range
    args = (sc.fmaps[0], sc.depths[0, :, :, 0], sc.intrs[0], sc.extrs[0])
    clouds = ops.cloud_build(*args, levels=[0, 1, 2], want_valid=True, want_tiles=True)
    plain = ops.cloud_build(*args, levels=[0, 1, 2], want_valid=True)
    for c in clouds:
        assert c.tile_z is not None and c.tile_z.shape == (S, V, (c.H + 7) // 8, (c.W + 7) // 8, 2)
        x4 = c.xyz4.view(S, V, c.H, c.W, 4).cpu().numpy().astype(np.float32)
        cam = c.cam.view(V, S, -1).cpu().numpy()
        z = np.empty((S, V, c.H, c.W), np.float32)
        for v in range(V):
            for s_ in range(S):
                r = cam[v, s_]
                # the kernel's fma chain in float32 is not reproducible with numpy products: compare with a tolerance
                z[s_, v] = (x4[s_, v, ..., 0].astype(np.float64) * r[30] + x4[s_, v, ..., 1].astype(np.float64) * r[31]
                            + x4[s_, v, ..., 2].astype(np.float64) * r[32] + r[35])
        z[(x4[..., 3] < 0) | np.isnan(z)] = np.nan
        TH, TW = (c.H + 7) // 8, (c.W + 7) // 8
        pad = np.full((S, V, TH * 8, TW * 8), np.nan, np.float32)
        pad[:, :, :c.H, :c.W] = z
        tiles = pad.reshape(S, V, TH, 8, TW, 8)
        with np.errstate(all="ignore"):
            import warnings
            with warnings.catch_warnings():
                warnings.simplefilter("ignore")
                lo, hi = np.nanmin(tiles, axis=(3, 5)), np.nanmax(tiles, axis=(3, 5))
        lo[np.isnan(lo)] = np.inf
        hi[np.isnan(hi)] = -np.inf
        got = c.tile_z.cpu().numpy()
        fin = np.isfinite(lo)
        assert (np.isfinite(got[..., 0]) == fin).all() and (np.isfinite(got[..., 1]) == fin).all()
        assert np.abs(got[..., 0][fin] - lo[fin]).max() < 1e-4 and np.abs(got[..., 1][fin] - hi[fin]).max() < 1e-4
        assert (got[..., 0][~fin] == np.inf).all() and (got[..., 1][~fin] == -np.inf).all()

    g = torch.Generator(device=DEV).manual_seed(91)
    coords = (torch.rand(S, 1200, 3, device=DEV, generator=g) - 0.5) * torch.tensor([12.0, 12.0, 4.0], device=DEV)
    E = sc.extrs[0, :, 0]
    centres = -(E[:, :, :3].transpose(1, 2) @ E[:, :, 3:]).squeeze(-1)
    coords[:, :600] = centres[torch.arange(600, device=DEV) % V] + 0.3 * torch.randn(S, 600, 3, device=DEV, generator=g)
    surf = clouds[0].xyz[:, torch.randint(0, clouds[0].xyz.shape[1], (300,), device=DEV, generator=g)]
    coords[:, 900:] = surf + 0.01 * torch.randn(S, 300, 3, device=DEV, generator=g)
    stats = torch.zeros(16, dtype=torch.int64, device=DEV)
    L = _lib.lib()
    for c, c0 in zip(clouds, plain):
        L.mvt_debug_set_stats(ctypes.c_void_p(stats.data_ptr()))
        stats.zero_()
        try:
            check_knn_raster(c, coords, 16)
            torch.cuda.synchronize()
        finally:
            L.mvt_debug_set_stats(None)
        tested, kept = int(stats[7]), int(stats[8])
        assert tested > 0 and kept < tested, (tested, kept)
        check_knn_raster(c, coords, 5, skip_invalid=True)
        d1, i1 = ops.knn(ops.Search.from_cloud(c, tiles=True), coords, 16)
        d0, i0 = ops.knn(ops.Search.from_cloud(c0), coords, 16)          # no tables: windows only
        assert c0.tile_z is None
        assert torch.equal(i0, i1) and torch.equal(d0, d1)
        # a single frame of the cloud (the reference's per-frame calls): the table is sliced with it
        d2_, i2_ = ops.knn(ops.Search.from_cloud(c, frame=1, tiles=True), coords[1:2].contiguous(), 16)
        assert torch.equal(i2_, i1[1:2]) and torch.equal(d2_, d1[1:2])
    # buffers reused in place for the next window: the old tables are stale and must be rebuilt before use
    sc2 = synth.make_scene(V, 80 * 4, 96 * 4, S=S, C=4, zero_frac=0.1, seed=92, device=DEV)
    again = ops.cloud_build(sc2.fmaps[0], sc2.depths[0, :, :, 0], sc2.intrs[0], sc2.extrs[0], levels=[0, 1, 2], want_valid=True,
                            reuse=clouds)
    assert all(not c.tiles_ready for c in again)
    for c in again:
        check_knn_raster(c, coords, 16)
        assert c.tiles_ready


def test_raster_knn_skewed_intrinsics_and_moving_cameras():
    sc = synth.make_scene(3, 40 * 4, 56 * 4, S=3, C=4, zero_frac=0.1, seed=50, device=DEV)
    g = torch.Generator(device=DEV).manual_seed(51)
    sc.intrs[..., 0, 1] = 7.5                                                      # skew
    sc.intrs[..., 0, 0] *= 1.2                                                     # fx != fy
    sc.extrs[..., 3] += 0.05 * torch.randn(sc.extrs[..., 3].shape, device=DEV, generator=g)   # per-frame cameras
    clouds = ops.cloud_build(sc.fmaps[0], sc.depths[0, :, :, 0], sc.intrs[0], sc.extrs[0], levels=[0, 2], want_valid=True)
    coords = _edge_queries(clouds[0], sc)
    for c in clouds:
        check_knn_raster(c, coords, 16)
    # non upper-triangular K: the raster search must notice and fall back to scanning whole rasters (still exact)
    sc.intrs[..., 1, 0] = 0.5
    c = ops.cloud_build(sc.fmaps[0], sc.depths[0, :, :, 0], sc.intrs[0], sc.extrs[0], levels=[1], want_valid=True)[0]
    check_knn_raster(c, coords[:, :40].contiguous(), 8)


def test_raster_knn_fewer_points_than_k_and_nan_depth():
    sc = synth.make_scene(2, 8 * 4, 8 * 4, S=1, C=4, zero_frac=0.0, seed=52, device=DEV)
    dp = sc.depths[0, :, :, 0].clone()
    dp[0, 0, 2, 3] = float("nan")
    dp[1, 0, :4] = float("inf")
    c = ops.cloud_build(sc.fmaps[0], dp, sc.intrs[0], sc.extrs[0], levels=[0, 3], want_valid=True)
    q = torch.randn(1, 30, 3, device=DEV)
    check_knn_raster(c[0], q, 16)
    dist, idx = ops.knn(c[1], q, 8)                # level 3 has 2 points only
    assert c[1].xyz.shape[1] == 2 and (idx[..., 2:] == -1).all() and (idx[..., 0] >= 0).all()


@pytest.mark.parametrize("scale", [1.3, 1.0 / 1.3, "shear"])
def test_raster_knn_non_rigid_extrinsics(scale):
    """the reference feeds similarity-scaled rotation blocks (apply_sim3_to_extrinsics writes R_inv / s,
    datasets/utils.py:403-415) and tolerates 1e-3 non-orthonormality.  A world ball of radius r is then a view-space
    ellipsoid inside a ball of radius sigma_max(R) * r: camera_prep stores that bound (cam[42]) and the pixel boxes use it."""
    sc = synth.make_scene(3, 45 * 4, 64 * 4, S=2, C=128, zero_frac=0.1, seed=70, device=DEV)
    if scale == "shear":
        A = torch.tensor([[1.0, 0.2, 0.0], [0.0, 0.9, 0.1], [0.05, 0.0, 1.15]], device=DEV)
        sc.extrs[..., :3, :3] = sc.extrs[..., :3, :3] @ A
    else:
        sc.extrs[..., :3, :3] *= scale
    clouds = ops.cloud_build(sc.fmaps[0], sc.depths[0, :, :, 0], sc.intrs[0], sc.extrs[0], levels=[0, 1, 3], want_valid=True)
    cam = clouds[0].cam.cpu().numpy()
    smax = np.linalg.svd(sc.extrs[0, :, 0, :3, :3].cpu().numpy().astype(np.float64), compute_uv=False).max(-1)
    assert (cam.reshape(3, 2, -1)[:, 0, 42] >= smax).all() and (cam.reshape(3, 2, -1)[:, 0, 42] <= smax * 1.75).all()
    coords = _edge_queries(clouds[0], sc)
    coords2, targets = synth.make_tracks(clouds[0].xyz[0], 200, 2, valid=clouds[0].valid[0], jitter=0.05)
    for c in clouds:
        check_knn_raster(c, coords, 16)
        check_knn_raster(c, coords2, 16)
        check_knn_raster(c, coords, 3, skip_invalid=True)
    check_corr(clouds[0].xyz, clouds[0].fvec, targets, coords2, 16, cloud=clouds[0])


def test_raster_knn_negative_depths_are_found():
    """the reference's knn (filter_invalid=False, mvtracker.py:816) considers EVERY unprojected point, also those
    with depth < 0 (behind their camera, on the backward half of the pixel's line).  cloud_build flags the frames that hold
    such points (mvt_level_out.neg_depth) and the raster search scans those frames exhaustively: still bit exact."""
    sc = synth.make_scene(2, 24 * 4, 32 * 4, S=3, C=128, zero_frac=0.1, seed=71, device=DEV)
    dp = sc.depths[0, :, :, 0].clone()
    dp[0, 1, 3:9, 5:20] *= -1.0                     # frame 1 of view 0: a patch of negative depths
    dp[1, 1, 0, 0] = -0.5
    clouds = ops.cloud_build(sc.fmaps[0], dp, sc.intrs[0], sc.extrs[0], levels=[0, 1], want_valid=True)
    assert clouds[0].neg.cpu().tolist() == [0, 1, 0]
    coords, targets = synth.make_tracks(clouds[0].xyz[0], 150, 3, valid=clouds[0].valid[0], jitter=0.05)
    neg_pts = clouds[0].xyz[1][(clouds[0].xyz4[1, :, 3] < 0)]
    coords[1, :40] = neg_pts[:40] + 0.01            # queries next to the points behind the camera
    for c in clouds:
        check_knn_raster(c, coords, 16)
        check_knn_raster(c, coords, 16, skip_invalid=True)      # "intended" filtering: depth <= 0 is not searchable
    check_corr(clouds[0].xyz, clouds[0].fvec, targets, coords, 16, cloud=clouds[0])
    # the next window in the same buffers has none: the flags are reset by the build
    again = ops.cloud_build(sc.fmaps[0], sc.depths[0, :, :, 0], sc.intrs[0], sc.extrs[0], levels=[0, 1], want_valid=True, reuse=clouds)
    assert again[0].neg.cpu().tolist() == [0, 0, 0]


def test_corr_warm_start_with_repeated_hint_indices():
    """the cap taken from idx_hint is the largest distance to k DISTINCT points; a hint that repeats an index
    (not something mvt_knn_corr ever writes, but the ABI accepts any int32) must not make the search inexact."""
    sc = synth.make_scene(3, 48 * 4, 64 * 4, S=2, C=128, zero_frac=0.1, seed=39, device=DEV)
    clouds = ops.cloud_build(sc.fmaps[0], sc.depths[0, :, :, 0], sc.intrs[0], sc.extrs[0], levels=[0, 2], want_valid=True)
    coords, targets = synth.make_tracks(clouds[0].xyz[0], 257, 2, valid=clouds[0].valid[0], jitter=0.03)
    lv = [(c, c.fvec, 16) for c in clouds]
    want, widx = ops.knn_corr(lv, targets, coords, return_idx=True)
    dup = widx.clone()
    dup[..., 1:] = dup[..., :1]                     # the nearest neighbour sixteen times: its distance is NOT a bound
    got, gidx = ops.knn_corr(lv, targets, coords, idx_hint=dup, return_idx=True)
    assert torch.equal(got, want) and torch.equal(gidx, widx)
    dup = widx.clone()
    dup[..., 15] = dup[..., 0]                      # one repeat
    got, gidx = ops.knn_corr(lv, targets, coords, idx_hint=dup, return_idx=True)
    assert torch.equal(got, want) and torch.equal(gidx, widx)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_corr_deferred_free_space_queries_identical(dtype):
    """Queries far from every surface (large pixel boxes) are handed by the fused kernel to knn_corr_hard_kernel (tile search,
    same gather / correlation code): outputs and neighbour lists must equal the one-kernel path bit for bit, and the list
    must really have been used."""
    sc = synth.make_scene(4, 80 * 4, 96 * 4, S=3, C=128, zero_frac=0.1, seed=95, device=DEV, dtype=dtype)
    clouds = ops.cloud_build(sc.fmaps[0], sc.depths[0, :, :, 0], sc.intrs[0], sc.extrs[0], levels=[0, 1, 2, 3], want_valid=True,
                             want_tiles=True)
    coords, targets = synth.make_tracks(clouds[0].xyz[0], 600, 3, valid=clouds[0].valid[0], dtype=dtype)
    g = torch.Generator(device=DEV).manual_seed(96)
    coords[:, :300] += 0.4 * torch.randn(3, 300, 3, device=DEV, generator=g)       # floating in free space
    E = sc.extrs[0, :, 0]
    centres = -(E[:, :, :3].transpose(1, 2) @ E[:, :, 3:]).squeeze(-1)
    coords[:, 300:340] = centres[torch.arange(40, device=DEV) % 4] + 0.2 * torch.randn(3, 40, 3, device=DEV, generator=g)
    lv = [(c, c.fvec, 16) for c in clouds]
    a, ia = ops.knn_corr(lv, targets, coords, return_idx=True, defer=False)
    s_, is_ = ops.knn_corr(lv, targets, coords, return_idx=True, defer="auto")      # whatever the policy picks: same bits
    assert torch.equal(is_, ia) and torch.equal(s_, a)
    b, ib = ops.knn_corr(lv, targets, coords, return_idx=True, defer=True)
    torch.cuda.synchronize()
    counters = ops._corr_ws[torch.cuda.current_device()][:8].view(torch.int32).cpu().tolist()
    assert counters[0] > 100 and counters[1] >= counters[0], counters      # deferred queries, all of them drawn
    assert torch.equal(ia, ib) and torch.equal(a, b)
    _, oi = oracle.knn(16, clouds[0].xyz.cpu().numpy(), coords.cpu().numpy())
    assert (ib[0].cpu().numpy() == oi).all()
    # count-only mode (what the adaptive policy runs while few queries are hard): one kernel, same bits, same count
    pol = ops.hard_query_policy(torch.device(DEV, torch.cuda.current_device()))
    pol.deferring, pol.event, pol.calls = False, None, 0      # calls = 0: the next launch is one the policy samples
    d_, id_ = ops.knn_corr(lv, targets, coords, return_idx=True, defer="auto")
    torch.cuda.synchronize()
    c2 = ops._corr_ws[torch.cuda.current_device()][:8].view(torch.int32).cpu().tolist()
    assert c2[0] == counters[0] and c2[1] == 0, (c2, counters)
    assert torch.equal(id_, ia) and torch.equal(d_, a)
    assert pol.mode() == 1 and pol.last_fraction > 0.05       # ... and the next call defers: many of these queries are hard
    # warm start + deferral (the cap travels with the deferred query)
    c, ic = ops.knn_corr(lv, targets, coords, return_idx=True, idx_hint=ia.clone(), defer=True)
    assert torch.equal(ic, ia) and torch.equal(c, a)
