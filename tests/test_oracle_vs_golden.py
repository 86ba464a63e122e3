"""CPU: pin the oracle (oracle/mvt_oracle.c) against golden vectors produced by the real reference."""
import numpy as np
import pytest

from oracle import oracle
from parity import assert_close, golden, knn_sets_equivalent


@pytest.mark.parametrize("name,g", golden("unproject"))
def test_unproject_matches_reference(name, g):
    mode = str(g["mode"])
    for lvl in g["levels"].tolist():
        xyz, fvec, valid = oracle.unproject(g["fmaps"], g["depths"], g["intrs"], g["extrs"], stride=int(g["stride"]),
                                            level=lvl, depth_interp_mode=mode, return_validity_mask=True)
        assert_close(xyz, g[f"xyz_l{lvl}"], 1e-4, f"{name} xyz l{lvl}")
        assert_close(fvec, g[f"fvec_l{lvl}"], 1e-5, f"{name} fvec l{lvl}")
        assert (valid == g[f"valid_l{lvl}"]).all()


@pytest.mark.parametrize("name,g", golden("knn"))
def test_knn_matches_reference(name, g):
    k = int(g["k"])
    dist, idx = oracle.knn(k, g["xyz"], g["query"])
    n_mis = knn_sets_equivalent(idx, g["idx"], g["xyz"], g["query"], name)
    # distances: reference cdist error grows with the scene offset (8.7e-3 at 10 units)
    tol = 2e-2 if name.endswith("_d") else 2e-3
    assert np.max(np.abs(np.sort(dist, -1) - np.sort(g["dists"], -1))) < tol
    # canonical order: ascending (d, idx)
    assert (np.diff(dist, axis=-1) >= 0).all()
    ties = np.diff(dist, axis=-1) == 0
    assert (np.diff(idx, axis=-1)[ties] > 0).all()
    print(name, "set mismatches adjudicated by f64:", n_mis)


@pytest.mark.parametrize("name,g", golden("corr"))
def test_corr_sample_matches_reference(name, g):
    k, G = int(g["k"]), int(g["groups"])
    out, idx, _ = oracle.corr_sample(k, g["xyz"], g["fvec"], g["targets"], g["coords"], G, bool(g["add_offset"]),
                                     bool(g["add_xyz"]), return_idx=True)
    ref = g["out"]
    assert out.shape == ref.shape
    # neighbour ORDER matters (mvtracker.py:374-384 consumes the k x (G+3) block positionally); rows whose
    # neighbour order differs from the reference only by ties/near-ties are compared as sorted multisets of rows
    a = out.reshape(-1, k, out.shape[-1])
    b = ref.reshape(-1, k, ref.shape[-1])
    exact_rows = 0
    for qa, qb in zip(a, b):
        if np.allclose(qa, qb, rtol=1e-4, atol=1e-5):
            exact_rows += 1
            continue
        sa = qa[np.lexsort(qa.T[::-1])]
        sb = qb[np.lexsort(qb.T[::-1])]
        assert np.allclose(sa, sb, rtol=1e-4, atol=1e-5), name
    assert exact_rows >= 0.95 * len(a), f"{name}: only {exact_rows}/{len(a)} queries match positionally"


@pytest.mark.parametrize("name,g", golden("input"))
def test_transformer_input_matches_reference(name, g):
    """Row f1: the oracle's assembly of the update transformer's input against the reference's own embedding functions
    joined as mvtracker.py:374-386 does (tests/golden/make_golden_input.py)."""
    E = int(g["flow_embed_dim"])
    x = oracle.transformer_input(g["coords"], g["fcorrs"], g["ffeats"], g["track_mask"], g["vis"], g["pos_embed"], g["time_embed"], E)
    assert x.shape == g["x"].shape
    assert np.abs(x - g["x"]).max() <= 5e-7                        # sin / cos columns: 1 ulp of values <= 1 (+ two adds)
    raw = oracle.transformer_input(g["coords"], g["fcorrs"], g["ffeats"], g["track_mask"], g["vis"], None, None, E)
    assert np.array_equal(raw[:, :, 3 * E:], g["x_noembed"][:, :, 3 * E:])       # copied columns: bit-exact
    assert np.abs(raw[:, :, :3 * E] - g["x_noembed"][:, :, :3 * E]).max() <= 1.2e-7


@pytest.mark.parametrize("name,g", golden("bwdcorr"))
def test_corr_backward_matches_reference_autograd(name, g):
    """Row f2: analytic gradients of corr_sample against torch autograd through the reference (make_golden_backward.py)."""
    k, G = int(g["k"]), int(g["groups"])
    out, idx, _ = oracle.corr_sample(k, g["xyz"], g["fvec"], g["targets"], g["coords"], G, True, bool(g["add_xyz"]), return_idx=True)
    assert_close(out, g["out"], 1e-4, name)
    gt, gf = oracle.corr_backward(g["fvec"], g["targets"], idx, g["w"], G)
    assert_close(gt, g["grad_targets"], 1e-5, f"{name} grad_targets")
    assert_close(gf, g["grad_fvec"], 1e-5, f"{name} grad_fvec")


@pytest.mark.parametrize("name,g", golden("bwdcloud"))
def test_unproject_backward_matches_reference_autograd(name, g):
    o = oracle.unproject_backward(g["w"], int(g["V"]), int(g["H"]), int(g["W"]), int(g["level"]))
    assert np.array_equal(o, g["grad_fmaps"])


@pytest.mark.parametrize("name,g", golden("predictor"))
def test_support_grid_matches_reference(name, g):
    """Row f3: the oracle's pixel lift against the reference's support-grid code (make_golden_predictor.py)."""
    V, T = g["depths"].shape[:2]
    pts = []
    for t in range(0, T, max(1, T // int(g["n_grids_per_view"]))):
        for v in range(V):
            w = oracle.lift_pixels(g["depths_resized"][v, t, 0], g["intrs_resized"][v, t], g["extrs"][v, t], g["pixel_xy"])
            pts.append(np.concatenate([np.full((len(w), 1), t, np.float32), w], 1))
    assert_close(np.concatenate(pts), g["support_points"], 1e-5, name)
