"""Row g: the drop-in inside the REAL model.  The reference's own MVTracker.forward
(mvtracker/models/core/mvtracker/mvtracker.py:412-732; imported from the unmodified copy that oracle/make_ref.py staged under
oracle/_ref, or from /root/reference) runs on the B200 with its stock seam and with mvtracker_b200.dropin.install();
traj_e / vis_e / feat_init are compared.  Random-init weights (no network), seed 72, BASELINE.json config 1 shape.

Two facts shape the bars (measured with oracle/model_harness.py, see DESIGN.md "Row g"):
  * the stock update head is initialised with std 1e-3 (cotracker2/blocks.py:443-451): an untrained model leaves every track
    where it started, so with the stock init the outputs are compared at the tolerances north_star names;
  * with a head that moves tracks (`lively`), the refinement loop is chaotic: the reference's OWN two kNN back ends --
    _knn_torch's |x|^2+|y|^2-2xy expansion vs exact differences -- already drift apart by centimetres, because one flipped
    16th neighbour changes where a track is asked next.  There the drop-in is held to the yardstick the reference sets for
    itself: its distance to the exact-distance arm must not exceed the stock arm's distance to that arm.
"""
import os
import sys

import pytest
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import ref_import  # noqa: E402

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not ref_import.available(),
                                                  reason="reference model not staged (python oracle/make_ref.py)")]

V, T, H, W, N, ITERS = 4, 24, 256, 256, 512, 4      # BASELINE.json configs[0]


@pytest.fixture(scope="module")
def seq():
    from oracle import model_harness as mh
    return mh.make_sequence(V, T, H, W, N, seed=72, device="cuda")


def _launches():
    from mvtracker_b200 import _lib
    return _lib.lib().mvt_launch_count()


def test_forward_stock_init_matches_reference(seq):
    from oracle import model_harness as mh
    model = mh.build_model(seed=72, device="cuda")
    ref_out, _ = mh.forward(model, seq, iters=ITERS, seam="stock")
    exact_out, _ = mh.forward(model, seq, iters=ITERS, seam="exact")
    n0 = _launches()
    new_out, _ = mh.forward(model, seq, iters=ITERS, seam="dropin")
    launched = _launches() - n0
    # 3 windows: camera_prep + cloud_build per window (+ feature-init build where queries start) and ONE fused knn_corr per
    # iteration -- not four; the per-level calls of the model hit the memos
    assert 3 * (2 + ITERS) <= launched <= 3 * (2 + ITERS) + 60, launched
    d, dx, yard = mh.compare(ref_out, new_out), mh.compare(exact_out, new_out), mh.compare(ref_out, exact_out)
    print("stock init: dropin vs stock", d, "| dropin vs exact-distance reference", dx, "| stock vs exact", yard)
    assert new_out["traj_e"].shape == (1, T, N, 3) and new_out["vis_e"].shape == (1, T, N)
    assert torch.isfinite(new_out["traj_e"]).all()
    # vs the reference as shipped: tracks whose initial feature row (k = 1 neighbour) is the same point
    assert d["traj_all"] <= 1e-4 * d["scale"], d
    assert d["vis"] <= 1e-3, d
    # the rows that differ are the reference's cdist expansion picking another point of a near-tie: against the
    # reference with exact distances they must (almost) all agree, and the stock reference must disagree with it just as often
    assert dx["feat_rows_equal"] >= 0.998 and dx["vis"] <= 1e-3 and dx["traj_all"] <= 1e-4 * dx["scale"], dx
    assert d["feat_rows_equal"] >= min(0.995, yard["feat_rows_equal"] - 0.004), (d, yard)


def test_forward_filter_invalid_depth_k8(seq):
    """demo.py:605-606 variant: corr_filter_invalid_depth=True, 8 neighbours (the reference runs its kNN on the compacted
    valid cloud and gathers from the uncompacted arrays, mvtracker.py:820-829; settings.filter_invalid_mode='reference')."""
    from oracle import model_harness as mh
    model = mh.build_model(seed=72, device="cuda", corr_filter_invalid_depth=True, corr_neighbors=8)
    ref_out, _ = mh.forward(model, seq, iters=2, seam="stock")
    new_out, _ = mh.forward(model, seq, iters=2, seam="dropin")
    d = mh.compare(ref_out, new_out)
    print("filter_invalid, k=8:", d)
    assert d["traj_all"] <= 1e-4 * d["scale"], d
    assert d["vis"] <= 1e-3 and d["feat_rows_equal"] >= 0.99, d


def test_forward_lively_head_tracks_like_exact_reference(seq):
    from oracle import model_harness as mh
    model = mh.build_model(seed=72, device="cuda", lively=0.002)
    stock, _ = mh.forward(model, seq, iters=ITERS, seam="stock")
    exact, _ = mh.forward(model, seq, iters=ITERS, seam="exact")
    ours, _ = mh.forward(model, seq, iters=ITERS, seam="dropin")
    start = seq.query_points[0, :, 1:]
    motion = (exact["traj_e"][0] - start).norm(dim=-1)                      # [T, N]

    def dev(a, b):
        e = (a["traj_e"][0] - b["traj_e"][0]).norm(dim=-1).amax(0)         # per track: worst frame
        return e.median().item(), e.quantile(0.95).item(), e.max().item()

    yard, mine = dev(stock, exact), dev(ours, exact)
    print(f"lively: motion mean {motion.mean().item():.4f} max {motion.max().item():.4f} | stock-vs-exact (median, p95, max) {yard} | "
          f"dropin-vs-exact {mine} | vis {mh.compare(ours, exact)['vis_all']:.2e} vs {mh.compare(stock, exact)['vis_all']:.2e}")
    assert motion.mean().item() > 1e-3, "the lively head does not move the tracks: the test would prove nothing"
    scale = float(exact["traj_e"].abs().max())
    # median track: at the fp32 tolerance; tails: no worse than the reference's own back ends differ (x2 for seed luck)
    assert mine[0] <= max(1e-4 * scale, 2 * yard[0]), (mine, yard)
    assert mine[1] <= max(1e-4 * scale, 2 * yard[1]), (mine, yard)
    assert mine[2] <= max(1e-3 * scale, 2 * yard[2]), (mine, yard)


def test_forward_under_inference_mode_and_bf16_autocast(seq):
    """inference tensors have no version counter; autocast feeds bf16 feature maps (demo.py:833-837)."""
    from oracle import model_harness as mh
    from oracle import ref_import as ri
    from mvtracker_b200 import dropin
    model = mh.build_model(seed=72, device="cuda")
    ref = ri.load()
    undo = dropin.install(ref.mvt)
    try:
        with torch.inference_mode():
            a = model(rgbs=seq.rgbs, depths=seq.depths, query_points=seq.query_points, intrs=seq.intrs, extrs=seq.extrs, iters=2)
    finally:
        undo()
    b, _ = mh.forward(model, seq, iters=2, seam="stock")
    d = mh.compare(b, a)
    print("inference_mode:", d)
    assert d["traj_all"] <= 1e-4 * d["scale"] and d["vis"] <= 1e-3 and d["feat_rows_equal"] >= 0.99, d
    s16, _ = mh.forward(model, seq, iters=2, seam="stock", autocast=torch.bfloat16)
    o16, _ = mh.forward(model, seq, iters=2, seam="dropin", autocast=torch.bfloat16)
    d = mh.compare(s16, o16)
    print("bf16 autocast:", d)
    # under autocast the reference's own unprojection einsums run in bf16 (model_utils.py:469-473: ~1 cm of error on the cloud
    # 4 m from the camera), so its neighbours -- and everything downstream of the features -- differ from the fp32 geometry of
    # the drop-in; the tracks (stock init: they stay where they start) and the shapes are what can be compared
    assert d["traj_all"] <= 2e-2 * d["scale"], d
    assert o16["vis_e"].shape == s16["vis_e"].shape and torch.isfinite(o16["vis_e"]).all() and torch.isfinite(o16["traj_e"]).all()


def test_forward_builds_every_frame_once(seq):
    """The per-frame cloud cache inside the real model: forward() walks T = 24 frames in three windows of 12 (hop 6,
    mvtracker.py:522-566); through the wrapped smart_cat the drop-in proves that the first 6 frames of windows 2 and 3 are
    the previous window's last 6 and builds 12 + 6 + 6 = 24 frames' clouds -- the reference (and a drop-in that could not
    prove the provenance) builds 36."""
    from oracle import model_harness as mh
    from mvtracker_b200 import dropin
    model = mh.build_model(seed=72, device="cuda")
    built = lambda: sum(s.frames_built for s in dropin._arenas.values())      # noqa: E731
    windows = lambda: sum(s.windows for s in dropin._arenas.values())        # noqa: E731
    b0, w0 = built(), windows()
    out, _ = mh.forward(model, seq, iters=2, seam="dropin")
    assert windows() - w0 == 3
    assert built() - b0 == T, f"{built() - b0} frames built for a {T}-frame sequence"
    assert torch.isfinite(out["traj_e"]).all()
