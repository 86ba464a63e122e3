import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np, torch
from mvtracker_b200 import ops, synth
from mvtracker_b200.dist import track_shard, view_range
from oracle import oracle
DEV = "cuda"
world = int(sys.argv[1]) if len(sys.argv) > 1 else 2
sc = synth.make_scene(4, 48 * 4, 64 * 4, S=3, C=128, zero_frac=0.1, seed=31, device=DEV)
fm, dp, K, E = sc.fmaps[0], sc.depths[0, :, :, 0], sc.intrs[0], sc.extrs[0]
clouds = ops.cloud_build(fm, dp, K, E, levels=range(4), want_valid=True)
coords, targets = synth.make_tracks(clouds[0].xyz[0], 203, 3, valid=clouds[0].valid[0], jitter=0.04)
coords[:, :5] += 2.5
fused, fidx = ops.knn_corr([(c, c.fvec, 16) for c in clouds], targets, coords, return_idx=True)
for r in range(world):
    v0, v1 = view_range(r, world, 4)
    cl = ops.cloud_build(fm[v0:v1], dp[v0:v1], K[v0:v1], E[v0:v1], levels=range(4), want_spacing=False)
    keys = ops.knn_keys(cl, [v0 * c.H * c.W for c in cl], coords, 16, n_shards=1)   # [1, L, S, M, k]
    for l, c in enumerate(cl):
        _, oi = oracle.knn(16, c.xyz.cpu().numpy(), coords.cpu().numpy())
        got = (keys[0, l] & 0xFFFFFFFF).cpu().numpy().astype(np.int64) - v0 * c.H * c.W
        miss = (keys[0, l] == ops.EMPTY_KEY).cpu().numpy()
        got[miss] = -1
        bad = (got != oi).any(-1)
        print(f"rank {r} level {l}: local knn_keys vs oracle on the local cloud: {bad.sum()} / {bad.size} queries differ")
        if bad.any():
            s_, m_ = np.argwhere(bad)[0]
            print("   first:", s_, m_, "got", got[s_, m_], "want", oi[s_, m_])
        # the same through mvt_knn
        _, ki = ops.knn(c, coords, 16, idx_dtype=torch.int32)
        print("   mvt_knn vs oracle:", int((ki.cpu().numpy() != oi).any(-1).sum()))

from test_gpu_shard import emulate
total, idx_all = emulate(sc, coords, targets, world, 16)
Ms = track_shard(world, 203)
merged = idx_all.permute(1, 2, 0, 3, 4).reshape(4, 3, world * Ms, 16)[:, :, :203]
bad = (merged != fidx).any(-1).cpu().numpy()
print("merged vs fused:", bad.sum(), "of", bad.size, "queries differ")
for l_, s_, m_ in np.argwhere(bad)[:6]:
    print(l_, s_, m_, "\n  merged", merged[l_, s_, m_].tolist(), "\n  fused ", fidx[l_, s_, m_].tolist())
print("by level:", bad.sum((1, 2)), "by frame:", bad.sum((0, 2)), "tracks:", np.unique(np.argwhere(bad)[:, 2])[:40])

# adjudicate with the oracle on the full cloud, and show the geometry of the first differing query
from mvtracker_b200 import _lib
for l_, s_, m_ in np.argwhere(bad)[:3]:
    c = clouds[l_]
    _, oi = oracle.knn(16, c.xyz[s_:s_ + 1].cpu().numpy(), coords[s_:s_ + 1, m_:m_ + 1].cpu().numpy())
    print("oracle ", oi[0, 0].tolist())
    grid = ops.build_grids([(c.xyz, None, None)], k_hint=16)[0]
    _, gi = ops.knn(grid, coords, 16, idx_dtype=torch.int32)
    print("grid   ", gi[s_, m_].tolist())
    _, ri = ops.knn(c, coords, 16, idx_dtype=torch.int32)
    print("raster ", ri[s_, m_].tolist())
    q = coords[s_, m_]
    only = sorted(set(oi[0, 0].tolist()) ^ set(fidx[l_, s_, m_].tolist()))
    print("q", q.tolist(), "H,W", c.H, c.W, "stride", c.stride)
    for i in only + oi[0, 0].tolist()[-3:]:
        p = c.xyz4[s_, i]
        v, pix = divmod(i, c.H * c.W)
        print(f"  point {i}: view {v} pixel (x={pix % c.W}, y={pix // c.W}) depth {p[3].item():.4f} d2 {float(((p[:3] - q) ** 2).sum()):.6f}")
    cam = c.cam.view(c.V, -1, 48)[:, s_]
    R, t = cam[:, 24:33].view(-1, 3, 3), cam[:, 33:36]
    qc = (R @ q) + t
    for v in range(c.V):
        X, Y, Z = qc[v].tolist()
        fx, sk, cx, fy, cy = cam[v, 36:41].tolist()
        fi = ((fx * X / Z + cx) + 0.5) / c.stride - 0.5
        fj = ((fy * Y / Z + cy) + 0.5) / c.stride - 0.5
        print(f"  view {v}: q_cam=({X:.3f},{Y:.3f},{Z:.3f}) raster coords ({fi:.3f},{fj:.3f})")
