"""Single-rank cost of the persistent sharded-iteration kernel (mvt_shard_iter, plain buffers) against the fused kernel and the
search-only kernel on the same window: what the formulation itself costs before any NVLink traffic.  python tools/diag_shard.py [workload] [frac]
frac: fraction of the tracks searched (emulates the per-rank share of phase A at N = 1 / frac ranks; phase B still gets every record)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mvtracker_b200 import ops, synth
dev = torch.device("cuda", 0)
w = synth.WORKLOADS[sys.argv[1] if len(sys.argv) > 1 else "panoptic"]
V, N = w["V"], w["N"]
sc = synth.make_scene(V, w["H"], w["W"], S=12, C=128, zero_frac=0.05, seed=72, device=dev)
clouds = ops.cloud_build(sc.fmaps[0], sc.depths[0, :, :, 0].contiguous(), sc.intrs[0].contiguous(), sc.extrs[0].contiguous(), levels=range(4), want_tiles=True)
lv = [(c, c.fvec, 16) for c in clouds]
def timed(fn, reps=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / reps
for n in (N, N // 2, N // 8):
    coords, targets = synth.make_tracks(clouds[0].xyz, n, 12, seed=73, valid=clouds[0].xyz4[..., 3] > 0)
    out = torch.empty(12, n, 256, device=dev)
    st = ops.ShardState.single(V, 4, 12, n, 16, dev)
    t_f = timed(lambda: ops.knn_corr(lv, targets, coords, out=out, defer=False))
    t_s = timed(lambda: ops.shard_iter(clouds, clouds, targets, coords, 16, st))
    t_k = sum(timed(lambda c=c: ops.knn(c, coords, 16, idx_dtype=torch.int32, want_dist=False)) for c in clouds)
    same = torch.equal(ops.shard_iter(clouds, clouds, targets, coords, 16, st), ops.knn_corr(lv, targets, coords, defer=False))
    print(f"{os.environ.get('MVT_LIB', 'default')[-24:]:24s} tracks {n:6d}: fused {t_f:.3f} ms | shard_iter (1 rank) {t_s:.3f} ms | search-only x4 levels {t_k:.3f} ms | equal {same} err {st.error()}", flush=True)
