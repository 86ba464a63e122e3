"""View-sharded kNN correlation (mvt_shard.cu): the three per-rank kernels, with the exchanges emulated in one process,
must reproduce the single-GPU fused kernel bit for bit."""
import numpy as np
import pytest
import torch

from mvtracker_b200 import ops, synth
from mvtracker_b200.dist import track_shard, view_range
from oracle import oracle

pytestmark = pytest.mark.gpu
DEV = "cuda"


def emulate(sc, coords, targets, world, k, levels=(0, 1, 2, 3), two_phase=True):
    fm, dp, K, E = sc.fmaps[0], sc.depths[0, :, :, 0], sc.intrs[0], sc.extrs[0]
    V = fm.shape[0]
    S, M, _ = targets.shape
    L, Ms = len(levels), track_shard(world, M)
    ranks = []
    for r in range(world):
        v0, v1 = view_range(r, world, V)
        clouds = ops.cloud_build(fm[v0:v1], dp[v0:v1], K[v0:v1], E[v0:v1], levels=levels, want_spacing=False) if v1 > v0 else []
        ranks.append((v0, v1, clouds))
    empty = torch.full((world, L, S, Ms, k), ops.EMPTY_KEY, dtype=torch.int64, device=DEV)

    flags = [torch.ones(world, L, S, Ms, dtype=torch.uint8, device=DEV) for _ in ranks]
    bufs = [empty.clone() for _ in ranks]

    def all_keys(caps=None, budget=0):
        for (v0, v1, clouds), fl, buf in zip(ranks, flags, bufs):
            if clouds:
                ops.knn_keys(clouds, [v0 * c.H * c.W for c in clouds], coords, k, n_shards=world, caps=caps, budget=budget,
                             exact_out=fl if budget else None, skip=fl if caps is not None else None, out=buf)
        return bufs

    caps = None
    if two_phase:      # phase A: bounded-effort searches -> merged k-th best = cap of the global k-th distance
        keys = all_keys(budget=two_phase)
        caps = torch.empty(world, L * S * Ms, dtype=torch.int64, device=DEV)
        for j in range(world):
            recv = torch.stack([keys[r][j] for r in range(world)]).contiguous()
            merged = torch.empty(L * S * Ms, k, dtype=torch.int64, device=DEV)
            ops.topk_merge(recv.view(world, L * S * Ms, k), keys_out=merged)
            caps[j] = merged[:, k - 1]
    keys = all_keys(caps=caps)
    idx_all = torch.empty(world, L, S, Ms, k, dtype=torch.int32, device=DEV)
    for j in range(world):      # all-to-all: rank j receives shard j of every rank; then merges
        recv = torch.stack([keys[r][j] for r in range(world)]).contiguous()
        idx_all[j] = ops.topk_merge(recv.view(world, L * S * Ms, k)).view(L, S, Ms, k)
    total = torch.zeros(S, M, L * k * 4, device=DEV)
    for v0, v1, clouds in ranks:
        if clouds:
            owned = [(c.xyz4, c.fvec, v0 * c.H * c.W, v1 * c.H * c.W) for c in clouds]
            total += ops.corr_from_idx(owned, idx_all, targets, coords, k)
    return total, idx_all


@pytest.mark.parametrize("world", [1, 2, 3, 5])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("two_phase", [0, 1, 150, 100000])     # 0: one uncapped phase; else the phase-A budget
def test_sharded_equals_fused(world, dtype, two_phase):
    sc = synth.make_scene(4, 48 * 4, 64 * 4, S=3, C=128, zero_frac=0.1, seed=31, device=DEV, dtype=dtype)
    clouds = ops.cloud_build(sc.fmaps[0], sc.depths[0, :, :, 0], sc.intrs[0], sc.extrs[0], levels=range(4), want_valid=True)
    coords, targets = synth.make_tracks(clouds[0].xyz[0], 203, 3, valid=clouds[0].valid[0], dtype=dtype, jitter=0.04)
    coords[:, :5] += 2.5
    fused, fidx = ops.knn_corr([(c, c.fvec, 16) for c in clouds], targets, coords, return_idx=True)
    total, idx_all = emulate(sc, coords, targets, world, 16, two_phase=two_phase)
    Ms = track_shard(world, 203)
    # idx_all [world, L, S, Ms, k] -> [L, S, M, k]
    merged = idx_all.permute(1, 2, 0, 3, 4).reshape(4, 3, world * Ms, 16)[:, :, :203]
    assert torch.equal(merged, fidx)
    assert torch.equal(total, fused)


def test_sharded_small_k_and_oracle():
    sc = synth.make_scene(3, 40 * 4, 56 * 4, S=2, C=128, zero_frac=0.3, seed=33, device=DEV)
    clouds = ops.cloud_build(sc.fmaps[0], sc.depths[0, :, :, 0], sc.intrs[0], sc.extrs[0], levels=[0, 2], want_valid=True)
    coords, targets = synth.make_tracks(clouds[0].xyz[0], 64, 2, valid=clouds[0].valid[0])
    total, idx_all = emulate(sc, coords, targets, 3, 8, levels=(0, 2))
    for i, c in enumerate(clouds):
        ref = oracle.corr_sample(8, c.xyz.cpu().numpy(), c.fvec.cpu().numpy(), targets.cpu().numpy(), coords.cpu().numpy())
        got = total[:, :, i * 32:(i + 1) * 32].reshape(2, 64, 8, 4).cpu().numpy()
        assert (got[..., 1:] == ref[..., 1:]).all()
        assert np.abs(got[..., 0] - ref[..., 0]).max() <= 1e-4 * max(1.0, np.abs(ref[..., 0]).max())


def test_shard_argument_errors():
    from mvtracker_b200._lib import MvtError
    sc = synth.make_scene(2, 32 * 4, 32 * 4, S=1, C=32, zero_frac=0.0, seed=34, device=DEV)
    c = ops.cloud_build(sc.fmaps[0], sc.depths[0, :, :, 0], sc.intrs[0], sc.extrs[0], levels=[0])[0]
    coords, targets = synth.make_tracks(c.xyz[0], 16, 1, C=32)
    idx = torch.zeros(1, 1, 1, 16, 16, dtype=torch.int32, device=DEV)
    with pytest.raises(MvtError, match="C == 128"):
        ops.corr_from_idx([(c.xyz4, c.fvec, 0, c.xyz4.shape[1])], idx, targets, coords, 16)


def _peers(tensors):
    """mvt_peers over plain device buffers of ONE process (no multicast address): every emulated rank's copy."""
    from mvtracker_b200._lib import Peers
    p = Peers()
    p.n = len(tensors)
    p.mc = None
    for r, t in enumerate(tensors):
        p.ptr[r] = t.data_ptr()
    return p


@pytest.mark.parametrize("world", [1, 2, 3, 8])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("split", [False, True])     # True: only correlation values are stored to the peers
def test_peer_store_kernels_equal_fused(world, dtype, split):
    """The "p2p" mode's three kernels (cloud_build_shared, knn_idx_shared, corr_from_idx_shared) with every rank's
    symmetric buffer emulated by a private buffer in this process: each rank's copy must equal the fused kernel's result."""
    V, S, M, k, L = 4, 3, 203, 16, 4
    sc = synth.make_scene(V, 48 * 4, 64 * 4, S=S, C=128, zero_frac=0.1, seed=41, device=DEV, dtype=dtype)
    fm, dp, K, E = sc.fmaps[0], sc.depths[0, :, :, 0], sc.intrs[0], sc.extrs[0]
    clouds = ops.cloud_build(fm, dp, K, E, levels=range(L), want_valid=True)
    coords, targets = synth.make_tracks(clouds[0].xyz[0], M, S, valid=clouds[0].valid[0], dtype=dtype, jitter=0.04)
    coords[:, :5] += 2.5
    fused, fidx = ops.knn_corr([(c, c.fvec, k) for c in clouds], targets, coords, return_idx=True)

    H, W = fm.shape[3], fm.shape[4]
    recs = [[torch.full((S, V * (H >> l) * (W >> l), 4), float("nan"), device=DEV) for l in range(L)] for _ in range(world)]
    local = []
    for r in range(world):
        v0, v1 = view_range(r, world, V)
        if v1 > v0:
            shared = {l: _peers([recs[j][l] for j in range(world)]) for l in range(L)}
            local.append(ops.cloud_build(fm[v0:v1], dp[v0:v1], K[v0:v1], E[v0:v1], levels=range(L), want_spacing=False,
                                         shared_records=shared, view_offset=v0, V_total=V))
        else:
            local.append([])
    for r in range(world):
        for l in range(L):
            assert torch.equal(recs[r][l], clouds[l].xyz4)          # every rank holds the whole window's records

    cam_all = ops.camera_prep(K, E)
    Ms = track_shard(world, M)
    idx_all = [torch.full((world, L, S, Ms, k), -7, dtype=torch.int32, device=DEV) for _ in range(world)]
    for r in range(world):
        full = [ops.full_cloud(l, recs[r][l], cam_all, V, H >> l, W >> l, 4.0 * (1 << l)) for l in range(L)]
        lo, hi = r * Ms, min((r + 1) * Ms, M)
        mine = coords[:, :1].repeat(1, Ms, 1).float().contiguous()
        if hi > lo:
            mine[:, :hi - lo] = coords[:, lo:hi]
        ops.knn_idx_shared(full, mine, k, r, world, _peers(idx_all))
    for r in range(world):
        got = idx_all[r].permute(1, 2, 0, 3, 4).reshape(L, S, world * Ms, k)[:, :, :M]
        assert torch.equal(got, fidx)

    outs = [torch.full((S, M, L * k * 4), float("nan"), device=DEV) for _ in range(world)]
    for r in range(world):
        v0, v1 = view_range(r, world, V)
        owned = [(c.xyz4, c.fvec, v0 * c.H * c.W, v1 * c.H * c.W) for c in local[r]]
        if owned:
            ops.corr_from_idx_shared(owned, idx_all[r], targets, coords, k, r, _peers(outs), outs[r].stride(1),
                                     replicas=recs[r] if split else None)
    for r in range(world):
        if split and not local[r]:
            continue            # a rank without views runs no kernel (dist.py's "p2p" mode needs a view on every rank)
        assert torch.equal(outs[r], fused)


@pytest.mark.parametrize("no_local", ["0", "1"])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_shard_iter_single_rank_equals_fused_kernel(dtype, no_local, monkeypatch):
    """mvt_shard_iter with one rank (plain buffers instead of symmetric memory): the persistent search -> inbox -> correlation
    kernel must reproduce the fused single-GPU kernel bit for bit, leave its inbox and counters empty for the next launch
    (two launches in a row, different queries), and handle queries with fewer than k searchable points.
    no_local = 1 (MVT_SHARD_NO_LOCAL): also the neighbours this rank owns travel through its inbox -- the path every record
    takes between ranks -- instead of being correlated by the searching warp."""
    monkeypatch.setenv("MVT_SHARD_NO_LOCAL", no_local)
    sc = synth.make_scene(3, 48 * 4, 64 * 4, S=3, C=128, zero_frac=0.2, seed=131, device=DEV, dtype=dtype)
    clouds = ops.cloud_build(sc.fmaps[0], sc.depths[0, :, :, 0], sc.intrs[0], sc.extrs[0], levels=[0, 1, 2, 3], want_tiles=True)
    M, k = 301, 16
    st = ops.ShardState.single(3, 4, 3, M, k, torch.device(DEV))
    for seed in (132, 133, 134):
        coords, targets = synth.make_tracks(clouds[0].xyz[0], M, 3, seed=seed, dtype=dtype, jitter=0.05)
        coords[:, :9] += 2.0                              # far from every surface: tile search inside the persistent kernel
        want = ops.knn_corr([(c, c.fvec, k) for c in clouds], targets, coords, defer=False)
        got = ops.shard_iter(clouds, clouds, targets, coords, k, st)
        torch.cuda.synchronize()
        assert not st.error()
        assert torch.equal(got, want), (seed, (got != want).any(-1).sum().item())
        assert int(st.inbox[0].abs().sum().item()) == 0 and int(st.local[:96].view(torch.int32).abs().sum().item()) == 0
        assert st.ctrl[0][:8].abs().sum().item() == 0 and int(st.ctrl[0][8].item()) == st.epoch
    # a coarse level with fewer points than k in a tiny raster: zero records come from the searching rank
    sc2 = synth.make_scene(2, 8 * 4, 8 * 4, S=2, C=128, zero_frac=0.0, seed=135, device=DEV, dtype=dtype)
    cl2 = ops.cloud_build(sc2.fmaps[0], sc2.depths[0, :, :, 0], sc2.intrs[0], sc2.extrs[0], levels=[0, 3], want_tiles=True)
    assert cl2[1].xyz.shape[1] == 2
    coords, targets = synth.make_tracks(cl2[0].xyz[0], 40, 2, seed=136, dtype=dtype)
    st2 = ops.ShardState.single(2, 2, 2, 40, 4, torch.device(DEV))
    got = ops.shard_iter(cl2, cl2, targets, coords, 4, st2)
    want = ops.knn_corr([(cl2[0], cl2[0].fvec, 4)], targets, coords, defer=False)
    torch.cuda.synchronize()
    assert torch.equal(got[:, :, :16], want) and (got[:, :, 16 + 8:] == 0).all() and not st2.error()


@pytest.mark.parametrize("prefetch", [False, True])
@pytest.mark.parametrize("world", [1, 2, 3, 8])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_frame_sharded_sequence_equals_fused_kernel(world, dtype, prefetch):
    """mode "frames" (dist.FrameShardedSequence): `world` emulated ranks, each owning the frames t % world == r with a cloud
    cache of its own, store through mvt_knn_corr_shared into one result block -- which must equal the single-GPU fused
    kernel on the freshly built window, for four sliding windows (cache reuse, wrap-around of the small arenas, per-frame
    cameras, ranks that add no frame to a window)."""
    from mvtracker_b200.dist import FrameShardedSequence
    V, Hs, Ws, S, hop, M, T = 3, 40, 56, 12, 6, 301, 30
    sc = synth.make_scene(V, Hs * 4, Ws * 4, S=T, C=128, zero_frac=0.1, seed=17, device=DEV, dtype=dtype)
    fm, dp, K = sc.fmaps[0], sc.depths[0, :, :, 0].contiguous(), sc.intrs[0].contiguous()
    E = sc.extrs[0].clone()
    E[:, :, :, 3] += 0.01 * torch.arange(T, device=DEV, dtype=torch.float32)[None, :, None]      # cameras differ per frame
    out = torch.empty(S, M, 256, device=DEV)
    most = (S + world - 1) // world
    # prefetch: window w + 1 is built on a side stream while window w's iterations run (needs 3 windows' worth of slots)
    seqs = [FrameShardedSequence(V, Hs, Ws, 128, dtype, DEV, S=S, hop=hop, rank=r, world=world, shared_out=out,
                                 slots=(3 if prefetch else 2) * most) for r in range(world)]
    n_win = 4

    def inputs(seq, w):
        ts = seq.new_frames()
        win = slice(hop * w, hop * w + S)
        return ts, (fm[:, ts].contiguous() if ts else None, dp[:, ts].contiguous() if ts else None, K[:, win], E[:, win])

    for w in range(n_win):
        t0 = hop * w
        win = slice(t0, t0 + S)
        full = ops.cloud_build(fm[:, win].contiguous(), dp[:, win].contiguous(), K[:, win].contiguous(), E[:, win].contiguous(),
                               levels=range(4), want_valid=True, want_tiles=True)
        coords, targets = synth.make_tracks(full[0].xyz, M, S, seed=40 + w, valid=full[0].valid, dtype=dtype)
        coords[:, :9] += 0.8                                      # free-space queries (the deferred path)
        want = ops.knn_corr([(c, c.fvec, 16) for c in full], targets, coords, defer=False)
        out.fill_(float("nan"))
        added = []
        for seq in seqs:
            if prefetch and w > 0:
                seq.push()                              # adopts what was prefetched during the previous window's iterations
            else:
                ts, args = inputs(seq, w)
                added += ts
                seq.push(*args)
        if not prefetch or w == 0:
            assert sorted(added) == list(range(t0, t0 + S) if w == 0 else range(t0 + S - hop, t0 + S))
        held = []
        for rep in range(2):
            for seq in seqs:
                assert seq.corr(targets, coords) is out
                if prefetch and rep == 0 and w + 1 < n_win:
                    ts, args = inputs(seq, w + 1)
                    held.append(args)                   # (inputs stay alive and untouched until the push)
                    seq.prefetch(*args)
            assert torch.equal(out, want), f"window {w}: {(out != want).any(-1).sum().item()} rows differ"
    assert sum(seq.frames_built for seq in seqs) == S + 3 * hop          # every frame was built exactly once, by its owner
