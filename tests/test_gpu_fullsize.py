"""Full-size checks on the windows BASELINE.json names and bench.py times -- DROID (3 views 720x1280, 2048 tracks), the
Panoptic-shaped window (8 views 1280x1600 -> 1 024 000 fused points per frame, 4096 tracks: the BENCH / SCALE workload) and
the DexYCB-shaped one (8 views 480x640, 16 384 tracks); stride 4, 12 frames, 4 levels, k = 16 -- through properties that do
not need an oracle pass over the whole problem:

* the oracle on a SAMPLE of queries against the whole 172 800-point frames (bit exact),
* two independent search structures -- the raster search and the hashed grid -- agree on every index of every query,
* canonical order of the lists, permutation equivariance over tracks, linearity of the correlation in the target features,
* the cloud against a plain torch fp32 restatement of the unprojection on the GPU (level 0 features are a pure transpose:
  bit exact).
"""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from mvtracker_b200 import ops, synth
from oracle import oracle

pytestmark = pytest.mark.gpu
DEV = "cuda"
S, K, L = 12, 16, 4
V = H = W = N = None        # set per workload by the fixture


@pytest.fixture(scope="module", params=["droid", "panoptic", "dexycb"])
def window(request):
    global V, H, W, N
    wl = synth.WORKLOADS[request.param]
    V, H, W, N = wl["V"], wl["H"], wl["W"], wl["N"]
    sc = synth.make_scene(V, H, W, S=S, C=128, zero_frac=0.05, seed=172, device=DEV)
    clouds = ops.cloud_build(sc.fmaps[0], sc.depths[0, :, :, 0], sc.intrs[0], sc.extrs[0], levels=range(L), want_valid=True)
    coords, targets = synth.make_tracks(clouds[0].xyz, N, S, seed=173, valid=clouds[0].valid)
    out, idx = ops.knn_corr([(c, c.fvec, K) for c in clouds], targets, coords, return_idx=True)
    torch.cuda.synchronize()
    yield sc, clouds, coords, targets, out, idx
    del sc, clouds, coords, targets, out, idx
    torch.cuda.empty_cache()


def test_cloud_against_torch_fp32(window):
    sc, clouds, *_ = window
    fm, dp, Kc, E = sc.fmaps[0], sc.depths[0, :, :, 0], sc.intrs[0], sc.extrs[0]       # [V,S,C,h,w], [V,S,h,w], [V,S,3,3], [V,S,3,4]
    h, w = fm.shape[-2:]
    f, d = fm.reshape(V * S, 128, h, w), dp.reshape(V * S, 1, h, w)
    Kinv = torch.inverse(Kc.double())
    E4 = torch.eye(4, device=DEV, dtype=torch.float64).repeat(V, S, 1, 1)
    E4[:, :, :3] = E.double()
    Einv = torch.inverse(E4)
    for c in clouds:
        l = c.level
        if l:
            f = F.avg_pool2d(f, 2, stride=2)
            d = F.interpolate(d, scale_factor=0.5, mode="nearest")
        hl, wl, st = h >> l, w >> l, 4.0 * (1 << l)
        # features: level 0 is a transpose (bit exact); pooled levels are chained 2x2 means in fp32
        want_f = f.reshape(V, S, 128, hl, wl).permute(1, 0, 3, 4, 2).reshape(S, V * hl * wl, 128)
        if l == 0:
            assert torch.equal(c.fvec, want_f)
        else:
            assert (c.fvec - want_f).abs().max().item() <= 1e-6 * max(1.0, want_f.abs().max().item())
        ys = (torch.arange(hl, device=DEV, dtype=torch.float64) + 0.5) * st - 0.5
        xs = (torch.arange(wl, device=DEV, dtype=torch.float64) + 0.5) * st - 0.5
        pix = torch.stack([xs[None, :].expand(hl, wl), ys[:, None].expand(hl, wl), torch.ones(hl, wl, device=DEV, dtype=torch.float64)], -1)
        cam = torch.einsum("vsij,hwj->vshwi", Kinv, pix) * d.reshape(V, S, hl, wl, 1).double()
        world = torch.einsum("vsij,vshwj->vshwi", Einv[:, :, :3, :3], cam) + Einv[:, :, None, None, :3, 3]
        want = world.permute(1, 0, 2, 3, 4).reshape(S, V * hl * wl, 3)
        ok = torch.isfinite(want).all(-1)
        err = (c.xyz.double() - want)[ok].abs().max().item()
        assert err <= 1e-4 * max(1.0, want[ok].abs().max().item()), (l, err)        # north_star: 1e-4 relative
        assert torch.equal(c.valid, (d.reshape(V, S, hl * wl) > 0).permute(1, 0, 2).reshape(S, V * hl * wl))


def test_sampled_queries_against_oracle(window):
    _, clouds, coords, targets, out, idx = window
    g = np.random.default_rng(7)
    for li in (0, 2, 3):
        c = clouds[li]
        for fr in (0, S - 1):
            pick = np.sort(g.choice(N, 48, replace=False))
            q = coords[fr, pick].cpu().numpy()[None]
            _, oi = oracle.knn(K, c.xyz[fr].cpu().numpy()[None], q)
            assert (idx[li, fr, pick].cpu().numpy() == oi[0]).all(), (li, fr)
            ref = oracle.corr_sample(K, c.xyz[fr].cpu().numpy()[None], c.fvec[fr].cpu().numpy()[None],
                                     targets[fr, pick].cpu().numpy()[None], q)
            got = out[fr, pick, li * K * 4:(li + 1) * K * 4].reshape(1, len(pick), K, 4).cpu().numpy()
            assert (got[..., 1:] == ref[..., 1:]).all()
            assert np.abs(got[..., 0] - ref[..., 0]).max() <= 1e-4 * max(1.0, np.abs(ref[..., 0]).max())


def test_raster_and_grid_searches_agree_everywhere(window):
    _, clouds, coords, _, _, idx = window
    grids = ops.build_grids([(c.xyz, None, c.spacing) for c in clouds], k_hint=K)
    for li, (c, gr) in enumerate(zip(clouds, grids)):
        dg, ig = ops.knn(gr, coords, K, idx_dtype=torch.int32)
        assert torch.equal(ig, idx[li]), f"level {li}: {(ig != idx[li]).any(-1).sum().item()} queries differ"
        dr, ir = ops.knn(c, coords, K, idx_dtype=torch.int32)           # search-only kernel (tile path on)
        assert torch.equal(ir, idx[li]) and torch.equal(dr, dg)
        # canonical order: distances ascending, ties by index; all indices distinct and in range
        assert (dr[..., 1:] >= dr[..., :-1]).all()
        # (dr is the square ROOT of the canonical d2: two different d2 may round to the same root, so a tie in dr binds the
        # index order only where the squared distances themselves are equal -- recomputed here in float64)
        tie = dr[..., 1:] == dr[..., :-1]
        if tie.any():
            fr, tr, j = torch.nonzero(tie, as_tuple=True)
            pa = c.xyz[fr, ir[fr, tr, j].long()].double() - coords[fr, tr].double()
            pb = c.xyz[fr, ir[fr, tr, j + 1].long()].double() - coords[fr, tr].double()
            da, db = (pa * pa).sum(-1), (pb * pb).sum(-1)
            real = da == db
            assert (ir[fr, tr, j + 1][real] > ir[fr, tr, j][real]).all()
            assert (db >= da * (1 - 1e-6)).all()
        assert ir.min().item() >= 0 and ir.max().item() < c.xyz.shape[1]
        assert (torch.sort(ir, -1).values.diff(dim=-1) > 0).all()


def test_track_permutation_and_target_linearity(window):
    _, clouds, coords, targets, out, _ = window
    lv = [(c, c.fvec, K) for c in clouds]
    perm = torch.randperm(N, device=DEV, generator=torch.Generator(device=DEV).manual_seed(5))
    out_p = ops.knn_corr(lv, targets[:, perm].contiguous(), coords[:, perm].contiguous())
    assert torch.equal(out_p, out[:, perm])
    t2 = torch.randn(targets.shape, device=DEV, generator=torch.Generator(device=DEV).manual_seed(6))
    o2 = ops.knn_corr(lv, t2, coords)
    o12 = ops.knn_corr(lv, (targets + t2).contiguous(), coords)
    a, b, c12 = [x.view(S, N, L * K, 4) for x in (out, o2, o12)]
    assert torch.equal(a[..., 1:], b[..., 1:]) and torch.equal(a[..., 1:], c12[..., 1:])      # offsets do not depend on features
    lin = (c12[..., 0] - (a[..., 0] + b[..., 0])).abs().max().item()
    assert lin <= 1e-4 * max(1.0, c12[..., 0].abs().max().item()), lin


@pytest.mark.parametrize("dist", ["static", "zero40"])
def test_other_track_distributions_sampled_against_oracle(window, dist):
    """The search cost -- and the code paths taken -- depend on where the tracks are.  `static`: the reference starts every
    window with the SAME 3-D point in all S frames (mvtracker.py:510), so tracks of moving objects float in free space in later
    frames; `zero40`: 40 % of the depth pixels invalid.  Sampled oracle check as above (the timed variants are
    bench.py's kernels.knn_corr_static / knn_corr_zero40)."""
    sc, clouds, coords, targets, _, _ = window
    if dist == "static":
        cl = clouds
        crd, _ = synth.make_tracks(cl[0].xyz[0], N, S, seed=176, valid=cl[0].valid[0])
    else:
        g = torch.Generator(device=DEV).manual_seed(177)
        dp = sc.depths[0, :, :, 0]
        dp40 = torch.where(torch.rand(dp.shape, device=DEV, generator=g) < 0.37, torch.zeros_like(dp), dp)
        cl = ops.cloud_build(sc.fmaps[0], dp40, sc.intrs[0], sc.extrs[0], levels=range(L), want_valid=True)
        crd, _ = synth.make_tracks(cl[0].xyz, N, S, seed=173, valid=cl[0].valid)
    out, idx = ops.knn_corr([(c, c.fvec, K) for c in cl], targets, crd, return_idx=True)
    rng = np.random.default_rng(8)
    for li in (0, 3):
        c = cl[li]
        for fr in (0, S - 1):
            pick = np.sort(rng.choice(N, 48, replace=False))
            q = crd[fr, pick].cpu().numpy()[None]
            _, oi = oracle.knn(K, c.xyz[fr].cpu().numpy()[None], q)
            assert (idx[li, fr, pick].cpu().numpy() == oi[0]).all(), (dist, li, fr)
            ref = oracle.corr_sample(K, c.xyz[fr].cpu().numpy()[None], c.fvec[fr].cpu().numpy()[None],
                                     targets[fr, pick].cpu().numpy()[None], q)
            got = out[fr, pick, li * K * 4:(li + 1) * K * 4].reshape(1, len(pick), K, 4).cpu().numpy()
            assert (got[..., 1:] == ref[..., 1:]).all()
            assert np.abs(got[..., 0] - ref[..., 0]).max() <= 1e-4 * max(1.0, np.abs(ref[..., 0]).max())
