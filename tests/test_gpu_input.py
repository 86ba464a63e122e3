"""GPU parity of mvt_track_input: the update transformer's input rows, against the golden vectors made
with the reference's own embedding functions and against the oracle on fresh inputs; fused with knn_corr's strided
output so that the [N, S, 581] buffer is written once."""
import numpy as np
import pytest
import torch

from mvtracker_b200 import ops, synth
from oracle import oracle
from parity import golden

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
# fp32 sin / cos of arguments up to ~1e3 * |flow|: CUDA's sinf / cosf are within 2 ulp of the correctly rounded value
TOL_TRIG = 5e-7


def _t(a):
    return torch.from_numpy(np.ascontiguousarray(a)).to(DEV)


@pytest.mark.parametrize("name,g", golden("input"))
def test_track_input_matches_reference_golden(name, g):
    E = int(g["flow_embed_dim"])
    S, N, Lc = g["fcorrs"].shape
    D = g["x"].shape[-1]
    x = torch.empty(N, S, D, device=DEV)
    ops.corr_columns(x, E, Lc).copy_(_t(g["fcorrs"]))
    got = ops.track_input(_t(g["coords"]), _t(g["ffeats"]), _t(g["track_mask"]), _t(g["vis"]), _t(g["pos_embed"]),
                          _t(g["time_embed"]), E, Lc, out=x, corr_in_out=True).cpu().numpy()
    assert np.abs(got - g["x"]).max() <= 2 * TOL_TRIG
    # without the embeddings every copied column is bit-exact and the correlation columns are left alone
    x.fill_(-3.0)
    raw = ops.track_input(_t(g["coords"]), _t(g["ffeats"]), _t(g["track_mask"]), _t(g["vis"]), None, None, E, Lc, out=x).cpu().numpy()
    c0 = 3 * E + 3
    assert (raw[:, :, c0:c0 + Lc] == -3.0).all()
    ref = g["x_noembed"]
    assert np.array_equal(raw[:, :, 3 * E:c0], ref[:, :, 3 * E:c0]) and np.array_equal(raw[:, :, c0 + Lc:], ref[:, :, c0 + Lc:])
    assert np.abs(raw[:, :, :3 * E] - ref[:, :, :3 * E]).max() <= TOL_TRIG


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_corr_and_input_written_once(dtype):
    """knn_corr writes its block into the [N, S, 581] buffer, track_input fills the rest and adds the embeddings: equal to
    the oracle's cat of (oracle embedding, the CUDA correlation block, features, masks) + pos + time."""
    S, N, K, E = 4, 203, 16, 64
    sc = synth.make_scene(3, 40 * 4, 56 * 4, S=S, C=128, zero_frac=0.1, seed=91, device=DEV, dtype=dtype)
    clouds = ops.cloud_build(sc.fmaps[0], sc.depths[0, :, :, 0], sc.intrs[0], sc.extrs[0], levels=range(4), want_valid=True)
    coords, targets = synth.make_tracks(clouds[0].xyz[0], N, S, valid=clouds[0].valid[0], dtype=dtype)
    lv = [(c, c.fvec, K) for c in clouds]
    dense = ops.knn_corr(lv, targets, coords)
    g = torch.Generator().manual_seed(3)
    D = 3 * E + 3 + 256 + 128 + 2
    pos, tim = torch.randn(N, D, generator=g).to(DEV), torch.randn(S, D, generator=g).to(DEV)
    mask = (torch.rand(S, N, generator=g) > 0.4).to(DEV)
    vis = torch.rand(S, N, generator=g).to(DEV)
    x = torch.full((N, S, D), float("nan"), device=DEV)
    ops.knn_corr(lv, targets, coords, out=ops.corr_columns(x, E, 256))
    ops.track_input(coords, targets, mask, vis, pos, tim, E, 256, out=x, corr_in_out=True)
    ref = oracle.transformer_input(coords.cpu().numpy(), dense.cpu().numpy(), targets.float().cpu().numpy(), mask.cpu().numpy(),
                                   vis.cpu().numpy(), pos.cpu().numpy(), tim.cpu().numpy(), E)
    got = x.cpu().numpy()
    assert np.isfinite(got).all()
    assert np.array_equal(got[:, :, 3 * E:], ref[:, :, 3 * E:])             # flow, corr, features, masks (+ embeddings): bit-exact
    assert np.abs(got[:, :, :3 * E] - ref[:, :, :3 * E]).max() <= 2 * TOL_TRIG


def test_track_input_full_size_properties():
    """Panoptic-size rows (4096 tracks x 12 frames x 581): frame 0 has zero flow (sin 0 = 0, cos 0 = 1 columns), the op is
    linear in the two embeddings, and a strided (padded) output buffer receives the same rows."""
    S, N, E, Lc, C = 12, 4096, 64, 256, 128
    g = torch.Generator().manual_seed(11)
    coords = (torch.randn(1, N, 3, generator=g) + 0.05 * torch.randn(S, N, 3, generator=g)).to(DEV)
    ff = torch.randn(S, N, C, generator=g).to(DEV)
    mask, vis = (torch.rand(S, N, generator=g) > 0.5).to(DEV), torch.rand(S, N, generator=g).to(DEV)
    D = 3 * E + 3 + Lc + C + 2
    pos, tim = torch.randn(N, D, generator=g).to(DEV), torch.randn(S, D, generator=g).to(DEV)
    x0 = torch.zeros(N, S, D, device=DEV)
    ops.track_input(coords, ff, mask, vis, None, None, E, Lc, out=x0)
    assert (x0[:, 0, 0:3 * E:2] == 0).all() and (x0[:, 0, 1:3 * E:2] == 1).all() and (x0[:, 0, 3 * E:3 * E + 3] == 0).all()
    assert torch.equal(x0[:, :, 3 * E + 3 + Lc:3 * E + 3 + Lc + C], ff.permute(1, 0, 2))
    assert torch.equal(x0[:, :, -2], mask.float().T) and torch.equal(x0[:, :, -1], vis.T)
    x1 = torch.zeros(N, S, D, device=DEV)
    ops.track_input(coords, ff, mask, vis, pos, tim, E, Lc, out=x1, corr_in_out=True)
    assert torch.equal(x1, (x0 + pos[:, None, :]) + tim[None, :, :])
    big = torch.zeros(N, S + 1, D + 3, device=DEV)
    view = big[:, :S, :D]
    ops.track_input(coords, ff, mask, vis, pos, tim, E, Lc, out=view, corr_in_out=True)
    assert torch.equal(view, x1) and (big[:, S] == 0).all() and (big[:, :, D:] == 0).all()


def test_track_input_rejects_cpu_tensors():
    from mvtracker_b200._lib import MvtError
    with pytest.raises(MvtError):
        ops.track_input(torch.zeros(2, 3, 3), torch.zeros(2, 3, 128), torch.zeros(2, 3), torch.zeros(2, 3), None, None)


def test_window_stream_delivers_padded_windows():
    """Row f4: pinned-host streaming of a sequence's windows (copy of window w+1 overlaps the work on window w); the last
    window repeats the final frame (mvtracker.py:561-566)."""
    from mvtracker_b200 import io as mio
    V, T, S = 2, 27, 12
    seq = {"fmaps": torch.randn(V, T, 4, 5, 6), "depths": torch.rand(V, T, 1, 5, 6)}
    seen = []
    for start, win in mio.WindowStream(seq, S=S, device=DEV):
        seen.append(start)
        for k, t in win.items():
            n = min(S, T - start)
            assert t.shape[1] == S and torch.equal(t[:, :n].cpu(), seq[k][:, start:start + n])
            if n < S:
                assert torch.equal(t[:, n:].cpu(), seq[k][:, T - 1:T].expand(-1, S - n, *seq[k].shape[2:]))
    assert seen == mio.window_starts(T, S) == [0, 6, 12, 18]


@pytest.mark.parametrize("name,g", golden("predictor"))
def test_predictor_preprocessing_matches_reference_golden(name, g):
    """Row f3: resize + intrinsics rescale + support grid (mvt_lift_pixels) against the reference's own code."""
    from mvtracker_b200 import predictor
    H, W = (int(v) for v in g["interp_shape"])
    rgbs, depths, intrs = predictor.resize_inputs(_t(g["rgbs"])[None], _t(g["depths"])[None], _t(g["intrs"])[None], (H, W))
    assert np.array_equal(rgbs[0].cpu().numpy(), g["rgbs_resized"]) and np.array_equal(depths[0].cpu().numpy(), g["depths_resized"])
    assert np.allclose(intrs[0].cpu().numpy(), g["intrs_resized"], rtol=1e-6)
    assert np.allclose(predictor.grid_pixels(int(g["grid_size"]), (H, W), DEV).cpu().numpy(), g["pixel_xy"], atol=1e-5)
    pts = predictor.support_grid_points(depths, intrs, _t(g["extrs"])[None], int(g["grid_size"]), int(g["n_grids_per_view"]))
    ref = g["support_points"]
    assert tuple(pts.shape) == (1,) + ref.shape
    assert np.array_equal(pts[0, :, 0].cpu().numpy(), ref[:, 0])
    err = np.abs(pts[0, :, 1:].cpu().numpy() - ref[:, 1:]).max()
    assert err <= 1e-4 * np.abs(ref[:, 1:]).max(), err


def test_lift_pixels_edges_vs_oracle():
    """Pixels outside the image (clamped corners, extrapolating weights), on integer positions, zero depth."""
    from mvtracker_b200 import ops as _ops
    sc = synth.make_scene(1, 20, 28, S=1, C=1, stride=1, zero_frac=0.2, seed=43, device=DEV)
    depth = sc.depths[0, 0, 0, 0]
    g = torch.Generator().manual_seed(4)
    pix = torch.cat([torch.rand(200, 2, generator=g) * torch.tensor([34.0, 26.0]) - 3.0,
                     torch.tensor([[0.0, 0.0], [27.0, 19.0], [27.5, 19.5], [-0.5, -0.5], [13.0, 7.0]])]).to(DEV)
    cam = _ops.camera_prep(sc.intrs[0, 0], sc.extrs[0, 0])[0]
    got = _ops.lift_pixels(depth, cam, pix).cpu().numpy()
    ref = oracle.lift_pixels(depth.cpu().numpy(), sc.intrs[0, 0, 0].cpu().numpy(), sc.extrs[0, 0, 0].cpu().numpy(), pix.cpu().numpy())
    assert np.abs(got - ref).max() <= 1e-4 * max(1.0, np.abs(ref).max())
    with_t = _ops.lift_pixels(depth, cam, pix, t=3.0).cpu().numpy()
    assert (with_t[:, 0] == 3.0).all() and np.array_equal(with_t[:, 1:], got)


@pytest.mark.parametrize("slots,dtype", [(18, torch.float32), (24, torch.float32), (30, torch.bfloat16)])
def test_window_sequence_frame_cache_equals_full_rebuild(slots, dtype):
    """Per-frame cloud cache across overlapping windows (the reference rebuilds all 12 frames of every window
    although 6 overlap, mvtracker.py:297-306 vs :549-555): WindowSequence.push builds only the 6 new frames into the arena;
    clouds, neighbour lists and correlation blocks of every window must equal a full rebuild bit for bit -- also across the
    wrap-around of a small arena, and with per-frame cameras (the kept frames keep THEIR cameras)."""
    from mvtracker_b200.hotpath import WindowClouds, WindowSequence
    V, Hs, Ws, T, S, hop = 3, 40, 56, 36, 12, 6
    sc = synth.make_scene(V, Hs * 4, Ws * 4, S=T, C=128, zero_frac=0.1, seed=101, device=DEV, dtype=dtype)
    g = torch.Generator(device=DEV).manual_seed(102)
    sc.extrs[..., 3] += 0.02 * torch.randn(sc.extrs[..., 3].shape, device=DEV, generator=g)        # moving cameras
    fm, dp, K, E = sc.fmaps[0], sc.depths[0], sc.intrs[0], sc.extrs[0]
    seq = WindowSequence(V, Hs, Ws, 128, dtype, torch.device(DEV), slots=slots, want_valid=True)
    for w in range((T - S) // hop + 1):
        f0 = w * hop
        new = slice(f0, f0 + S) if w == 0 else slice(f0 + S - hop, f0 + S)
        win = seq.push(fm[:, new], dp[:, new], K[:, f0:f0 + S], E[:, f0:f0 + S])
        ref = WindowClouds.build(fm[:, f0:f0 + S], dp[:, f0:f0 + S], K[:, f0:f0 + S], E[:, f0:f0 + S])
        for a, b in zip(win.clouds, ref.clouds):
            assert torch.equal(a.xyz, b.xyz) and torch.equal(a.fvec, b.fvec) and torch.equal(a.xyz4, b.xyz4), (w, a.level)
            assert torch.equal(a.tile_z, b.tile_z) and a.neg.sum().item() == 0
        coords, targets = synth.make_tracks(ref.clouds[0].xyz, 200, S, seed=103 + w, dtype=dtype, valid=ref.clouds[0].xyz4[..., 3] > 0)
        coords[:, :20] += 0.4                                       # some free-space queries (deferred to the tile search)
        assert torch.equal(win.corr(targets, coords), ref.corr(targets, coords)), w
        assert torch.equal(win.feature_init(3, coords[3, :50]), ref.feature_init(3, coords[3, :50]))
    assert seq.frames_built == S + hop * ((T - S) // hop)


def test_torch_library_ops_equal_the_ctypes_path_and_trace_without_graph_breaks():
    """torch.ops.mvt.* give the tensors of mvtracker_b200.ops bit for bit, and a function around them compiles with
    fullgraph=True (no graph break at the custom ops; the "eager" backend keeps the test independent of a code generator)."""
    import mvtracker_b200.torch_ops  # noqa: F401
    V, Hs, Ws, S, N, K = 3, 40, 56, 3, 157, 16
    sc = synth.make_scene(V, Hs * 4, Ws * 4, S=S, C=128, zero_frac=0.1, seed=81, device=DEV)
    fm, dp, Ki, E = sc.fmaps[0], sc.depths[0, :, :, 0].contiguous(), sc.intrs[0].contiguous(), sc.extrs[0].contiguous()
    clouds = ops.cloud_build(fm, dp, Ki, E, levels=range(4))
    outs = torch.ops.mvt.cloud_build(fm, dp, Ki, E, 4.0, [0, 1, 2, 3], False)
    for i, c in enumerate(clouds):
        assert torch.equal(outs[3 * i], c.xyz) and torch.equal(outs[3 * i + 1], c.fvec) and torch.equal(outs[3 * i + 2], c.xyz4)
    cam = outs[-1]
    coords, targets = synth.make_tracks(clouds[0].xyz[0], N, S)
    want = ops.knn_corr([(c, c.fvec, K) for c in clouds], targets, coords)

    def f(x4, fv, cam, t, q):
        corr = torch.ops.mvt.knn_corr(x4, fv, cam, t * 1.0, q + 0.0, K, V, Hs, Ws, 4.0)
        d, i = torch.ops.mvt.knn(x4[0], cam, q, 4, V, Hs, Ws, 4.0)
        return corr * 2.0, d, i

    x4, fv = list(outs[2::3][:4]), list(outs[1::3][:4])
    got, d, i = torch.compile(f, backend="eager", fullgraph=True)(x4, fv, cam, targets, coords)
    assert torch.equal(got, want * 2.0)
    d0, i0 = ops.knn(clouds[0], coords, 4)
    assert torch.equal(d, d0) and torch.equal(i, i0)


def test_refinement_iteration_in_a_cuda_graph():
    """the C ABI only enqueues on the current stream, so one refinement iteration's hot path -- the fused
    kNN correlation of all levels written into the transformer-input buffer + mvt_track_input -- is captured in a
    torch.cuda.CUDAGraph and replayed with new track positions / features; every replay must equal the eager result.
    (The reference's loop has a host sync per iteration, `torch.isnan(coords).any()` at mvtracker.py:401; INTEGRATION.md
    shows how a captured loop defers that check.)"""
    from mvtracker_b200.hotpath import WindowClouds
    V, Hs, Ws, S, N = 3, 40, 56, 4, 200
    sc = synth.make_scene(V, Hs * 4, Ws * 4, S=S, C=128, zero_frac=0.1, seed=83, device=DEV)
    win = WindowClouds.build(sc.fmaps, sc.depths, sc.intrs, sc.extrs)
    D = 581
    g = torch.Generator(device=DEV).manual_seed(84)
    pos, tim = torch.randn(N, D, device=DEV, generator=g), torch.randn(S, D, device=DEV, generator=g)
    mask, vis = torch.ones(S, N, device=DEV), torch.rand(S, N, device=DEV, generator=g)
    # static input / output buffers of the captured graph
    coords_b, feats_b = torch.empty(S, N, 3, device=DEV), torch.empty(S, N, 128, device=DEV)
    x_b = torch.empty(N, S, D, device=DEV)
    sets = []
    for i in range(4):
        c, t = synth.make_tracks(win.clouds[0].xyz[0], N, S, seed=85 + i, jitter=0.02 + 0.1 * i)      # incl. free-space queries
        sets.append((c, t))
    coords_b.copy_(sets[0][0]); feats_b.copy_(sets[0][1])
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):                       # warm-up on the capture stream (allocations, workspace)
        win.transformer_input(feats_b, coords_b, mask, vis, pos, tim, out=x_b)
    torch.cuda.current_stream().wait_stream(side)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        win.transformer_input(feats_b, coords_b, mask, vis, pos, tim, out=x_b)
    for c, t in sets[1:] + sets[:1]:
        coords_b.copy_(c); feats_b.copy_(t)
        graph.replay()
        torch.cuda.synchronize()
        want = win.transformer_input(t, c, mask, vis, pos, tim)
        assert torch.equal(x_b, want)
