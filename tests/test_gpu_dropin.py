"""GPU: the host mirror of the reference seam (mvtracker_b200.dropin) driven exactly the way
MVTracker.forward / forward_iteration drive the reference objects (mvtracker.py:296-322, :350-376, :606-645)."""
import numpy as np
import pytest
import torch

from mvtracker_b200 import dropin, synth
from oracle import oracle
from parity import assert_close, golden, knn_sets_equivalent

pytestmark = pytest.mark.gpu
DEV = "cuda"


def t(a):
    return torch.as_tensor(np.ascontiguousarray(a)).to(DEV)


@pytest.fixture(autouse=True)
def _fresh():
    dropin.clear_caches()
    dropin.settings.pyramid_cache = True
    dropin.settings.fuse_levels = True
    dropin.settings.filter_invalid_mode = "reference"
    dropin.settings.search = "raster"
    yield
    dropin.clear_caches()


@pytest.mark.parametrize("cache", [True, False])
@pytest.mark.parametrize("name,g", golden("unproject"))
def test_init_pointcloud_golden(name, g, cache):
    dropin.settings.pyramid_cache = cache
    fm, dp = t(g["fmaps"])[None], t(g["depths"])[None, :, :, None]
    K, E = t(g["intrs"])[None], t(g["extrs"])[None]
    for lvl in g["levels"].tolist():
        xyz, fvec, valid = dropin.init_pointcloud_from_rgbd(fm, dp, K, E, stride=int(g["stride"]), level=lvl,
                                                            depth_interp_mode=str(g["mode"]), return_validity_mask=True)
        assert xyz.shape == g[f"xyz_l{lvl}"].shape and fvec.shape == g[f"fvec_l{lvl}"].shape
        assert valid.dtype == torch.bool
        assert_close(xyz.cpu().numpy(), g[f"xyz_l{lvl}"], 1e-4, f"{name} xyz l{lvl}")
        assert_close(fvec.cpu().numpy(), g[f"fvec_l{lvl}"], 1e-5, f"{name} fvec l{lvl}")
        assert (valid.cpu().numpy() == g[f"valid_l{lvl}"]).all()
        pair = dropin.init_pointcloud_from_rgbd(fm, dp, K, E, stride=int(g["stride"]), level=lvl,
                                                depth_interp_mode=str(g["mode"]))
        assert len(pair) == 2


def test_pyramid_cache_invalidation():
    sc = synth.make_scene(2, 32 * 4, 32 * 4, S=2, C=8, seed=1, device=DEV)
    a0, _ = dropin.init_pointcloud_from_rgbd(sc.fmaps, sc.depths, sc.intrs, sc.extrs, level=0)
    a1, _ = dropin.init_pointcloud_from_rgbd(sc.fmaps, sc.depths, sc.intrs, sc.extrs, level=1)
    again, _ = dropin.init_pointcloud_from_rgbd(sc.fmaps, sc.depths, sc.intrs, sc.extrs, level=0)
    assert again.data_ptr() == a0.data_ptr()                     # memoised
    sc.depths.mul_(2.0)                                          # in-place edit bumps _version -> rebuild
    b0, _ = dropin.init_pointcloud_from_rgbd(sc.fmaps, sc.depths, sc.intrs, sc.extrs, level=0)
    assert b0.data_ptr() != a0.data_ptr() and not torch.equal(b0, a0)


@pytest.mark.parametrize("name,g", golden("knn"))
def test_knn_golden(name, g):
    k = int(g["k"])
    dists, idx = dropin.knn(k, t(g["xyz"]), t(g["query"]))
    assert idx.dtype == torch.int64 and dists.shape == g["dists"].shape
    od, oi = oracle.knn(k, g["xyz"], g["query"])
    assert (idx.cpu().numpy() == oi).all() and (dists.cpu().numpy() == od).all()
    knn_sets_equivalent(idx.cpu().numpy(), g["idx"], g["xyz"], g["query"], name)


@pytest.mark.parametrize("name,g", golden("corr"))
def test_corr_block_golden(name, g):
    k, G = int(g["k"]), int(g["groups"])
    blk = dropin.PointcloudCorrBlock(k=k, groups=G, xyz=t(g["xyz"]), fvec=t(g["fvec"]),
                                     corr_add_neighbor_offset=bool(g["add_offset"]), corr_add_neighbor_xyz=bool(g["add_xyz"]))
    out = blk.corr_sample(t(g["targets"]), t(g["coords"]))
    assert out.shape == g["out"].shape
    o_out = oracle.corr_sample(k, g["xyz"], g["fvec"], g["targets"], g["coords"], G, bool(g["add_offset"]), bool(g["add_xyz"]))
    assert_close(out.cpu().numpy(), o_out, 1e-4, name)
    same = np.isclose(out.cpu().numpy(), g["out"], rtol=1e-4, atol=1e-5).reshape(-1, k * out.shape[-1]).all(-1)
    assert same.mean() >= 0.95


@pytest.mark.parametrize("search", ["raster", "grid"])
@pytest.mark.parametrize("fuse", [True, False])
def test_forward_iteration_call_pattern(fuse, search):
    """4 levels built level by level, 4 blocks, then iters x levels corr_sample calls, concatenated as at mvtracker.py:374."""
    dropin.settings.fuse_levels = fuse
    dropin.settings.search = search
    V, S, N, C, k = 3, 12, 200, 128, 16
    sc = synth.make_scene(V, 48 * 4, 64 * 4, S=S, C=C, zero_frac=0.1, seed=31, device=DEV)
    blocks, clouds = [], []
    for lvl in range(4):
        xyz, fvec = dropin.init_pointcloud_from_rgbd(sc.fmaps, sc.depths, sc.intrs, sc.extrs, stride=4, level=lvl)
        clouds.append((xyz, fvec))
        blocks.append(dropin.PointcloudCorrBlock(k=k, groups=1, xyz=xyz, fvec=fvec, corr_add_neighbor_offset=True,
                                                 corr_add_neighbor_xyz=False))
    coords, ffeats = synth.make_tracks(clouds[0][0][0], N, S, C=C)
    for it in range(2):
        coords = coords + 0.01 * (it + 1)
        ffeats = ffeats * 0.9
        fcorrs = [blocks[l].corr_sample(ffeats.reshape(S, N, C), coords.reshape(S, N, 3)).reshape(1, S, N, -1) for l in range(4)]
        fcorrs = torch.cat(fcorrs, dim=-1)
        assert fcorrs.shape == (1, S, N, 256)
        for l in range(4):
            o = oracle.corr_sample(k, clouds[l][0].cpu().numpy(), clouds[l][1].cpu().numpy(), ffeats.cpu().numpy(),
                                   coords.cpu().numpy())
            assert_close(fcorrs[0, :, :, l * 64:(l + 1) * 64].cpu().numpy(), o.reshape(S, N, 64), 1e-4, f"it {it} level {l}")


@pytest.mark.parametrize("search", ["raster", "grid"])
def test_feature_init_call_pattern(search):
    """k = 1 kNN of query points against one frame's fused level-0 cloud, then a row gather (mvtracker.py:627-643)."""
    dropin.settings.search = search
    sc = synth.make_scene(3, 40 * 4, 64 * 4, S=6, C=128, zero_frac=0.1, seed=32, device=DEV)
    xyz, fvec = dropin.init_pointcloud_from_rgbd(sc.fmaps, sc.depths, sc.intrs, sc.extrs, stride=4)
    q, _ = synth.make_tracks(xyz[0], 50, 1)
    if search == "raster":
        assert dropin._find_raster(xyz[3][None]) is not None and dropin._find_raster(xyz[3][None].clone()) is None
    for f in range(2):
        dists, idx = dropin.knn(1, xyz[f][None], q[0][None])
        feat = fvec[f][idx[0, :, 0]]
        od, oi = oracle.knn(1, xyz[f][None].cpu().numpy(), q[0][None].cpu().numpy())
        assert (idx.cpu().numpy() == oi).all()
        assert torch.equal(feat, fvec[f][torch.as_tensor(oi[0, :, 0]).long().to(DEV)])


@pytest.mark.parametrize("mode", ["reference", "intended"])
def test_filter_invalid(mode):
    dropin.settings.filter_invalid_mode = mode
    sc = synth.make_scene(2, 32 * 4, 48 * 4, S=3, C=32, zero_frac=0.3, seed=33, device=DEV)
    xyz, fvec, valid = dropin.init_pointcloud_from_rgbd(sc.fmaps, sc.depths, sc.intrs, sc.extrs, return_validity_mask=True)
    coords, targets = synth.make_tracks(xyz[0], 64, 3, C=32, valid=valid[0])
    blk = dropin.PointcloudCorrBlock(k=8, groups=1, xyz=xyz, fvec=fvec, corr_add_neighbor_offset=True,
                                     corr_add_neighbor_xyz=False, filter_invalid=True, valid=valid)
    out = blk.corr_sample(targets, coords).cpu().numpy()
    x, f, v = xyz.cpu().numpy(), fvec.cpu().numpy(), valid.cpu().numpy()
    for b in range(3):
        if mode == "reference":
            # mvtracker.py:820-829: kNN on the compacted cloud, indices then applied to the UNcompacted arrays
            _, ci = oracle.knn(8, x[b][v[b]][None], coords[b][None].cpu().numpy())
            idx = ci
        else:
            _, idx = oracle.knn(8, x[b][None], coords[b][None].cpu().numpy(), mask=v[b][None])
        o = oracle.corr_from_idx(x[b][None], f[b][None], targets[b][None].cpu().numpy(), coords[b][None].cpu().numpy(), idx)
        assert_close(out[b][None], o, 1e-4, f"{mode} frame {b}")


def test_pointops_shaped_knn_query_with_unequal_segments():
    """mvtracker_b200.pointops_compat.knn_query(k, xyz, offset, new_xyz, new_offset): the signature of the reference's native
    precedent (pointops.knn_query, mvtracker.py:55-65) -- flattened points, cumulative offsets, GLOBAL int32 indices -- with
    segments of DIFFERENT sizes, against the canonical oracle per segment; and the reference's own un-offset arithmetic
    (mvtracker.py:68-69) on an equal-sized call."""
    from mvtracker_b200.pointops_compat import knn_query
    g = torch.Generator(device=DEV).manual_seed(11)
    sizes, qsizes, k = [5000, 37, 1200, 9], [300, 11, 1, 25], 16
    xyz = torch.rand(sum(sizes), 3, device=DEV, generator=g) * 3
    new = torch.rand(sum(qsizes), 3, device=DEV, generator=g) * 3
    off = torch.tensor(np.cumsum(sizes), device=DEV, dtype=torch.int32)
    noff = torch.tensor(np.cumsum(qsizes), device=DEV, dtype=torch.int32)
    idx, dist = knn_query(k, xyz, off, new, noff)
    assert idx.dtype == torch.int32 and idx.shape == (sum(qsizes), k) and dist.shape == idx.shape
    a = qa = 0
    for n, m in zip(sizes, qsizes):
        od, oi = oracle.knn(min(k, 32), xyz[a:a + n][None].cpu().numpy(), new[qa:qa + m][None].cpu().numpy())
        got = idx[qa:qa + m].cpu().numpy()
        want = np.where(oi[0] >= 0, oi[0] + a, -1)
        assert (got == want).all(), (n, m)
        assert (dist[qa:qa + m].cpu().numpy()[want >= 0] == od[0][want >= 0]).all()
        assert (got[:, n:] == -1).all()                    # fewer than k points in the segment
        a, qa = a + n, qa + m
    # the call the reference makes: B equal segments, then idx - b*N  (mvtracker.py:60-69)
    B, N, M = 3, 700, 40
    ref, qry = torch.rand(B, N, 3, device=DEV, generator=g), torch.rand(B, M, 3, device=DEV, generator=g)
    offset = torch.arange(1, B + 1, device=DEV) * N
    new_offset = torch.arange(1, B + 1, device=DEV) * M
    idx, dist = knn_query(k, ref.view(B * N, 3), offset, qry.view(B * M, 3), new_offset)
    idx = idx.view(B, M, k) - (torch.arange(B, device=DEV).view(B, 1, 1) * N)
    od, oi = oracle.knn(k, ref.cpu().numpy(), qry.cpu().numpy())
    assert (idx.cpu().numpy() == oi).all() and (dist.view(B, M, k).cpu().numpy() == od).all()
