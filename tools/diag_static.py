"""Per-distribution timing of one fused knn_corr launch on the panoptic window: surface-following / static-init / 40 % invalid
depth, with and without deferral of the large-box queries; prints the deferred-query counts.  python tools/diag_static.py [workload]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mvtracker_b200 import ops, synth
dev = torch.device("cuda", 0)
w = synth.WORKLOADS[sys.argv[1] if len(sys.argv) > 1 else "panoptic"]
V, N = w["V"], w["N"]
dt = torch.bfloat16 if (len(sys.argv) > 2 and sys.argv[2] == "bf16") else torch.float32
sc = synth.make_scene(V, w["H"], w["W"], S=12, C=128, zero_frac=0.05, seed=72, device=dev, dtype=dt)
clouds = ops.cloud_build(sc.fmaps[0], sc.depths[0, :, :, 0].contiguous(), sc.intrs[0].contiguous(), sc.extrs[0].contiguous(), levels=range(4), want_valid=True, want_tiles=True)
lv = [(c, c.fvec, 16) for c in clouds]
surf, targets = synth.make_tracks(clouds[0].xyz, N, 12, seed=73, valid=clouds[0].valid, dtype=dt)
stat, _ = synth.make_tracks(clouds[0].xyz[0], N, 12, seed=76, valid=clouds[0].valid[0])
out = torch.empty(12, N, 256, device=dev)
def timed(coords, defer, reps=10):
    for _ in range(3): ops.knn_corr(lv, targets, coords, out=out, defer=defer)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps): ops.knn_corr(lv, targets, coords, out=out, defer=defer)
    b.record(); torch.cuda.synchronize()
    cnt = ops._corr_ws[0][:8].view(torch.int32).cpu().tolist()[0] if (defer and 0 in ops._corr_ws) else 0
    return a.elapsed_time(b) / reps, cnt
for name, c in (("surface", surf), ("static", stat)):
    t0, _ = timed(c, False)
    t1, n = timed(c, True)
    print(f"{os.environ.get('MVT_LIB', 'default'):40s} {name:8s} one kernel {t0:.3f} ms | deferral {t1:.3f} ms, {n} of {4 * 12 * N} queries deferred", flush=True)
