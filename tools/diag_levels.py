"""Diagnostics (GPU): where the fused knn_corr launch spends its time -- per pyramid level, fused kernel vs search only
(mvt_knn: same raster search, no gather).    python tools/diag_levels.py [--workload panoptic]"""
import argparse, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mvtracker_b200 import ops, synth

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="panoptic")
a = ap.parse_args()
w = synth.WORKLOADS[a.workload]
dev = "cuda"
sc = synth.make_scene(w["V"], w["H"], w["W"], S=12, C=128, zero_frac=0.05, seed=72, device=dev)
clouds = ops.cloud_build(sc.fmaps[0], sc.depths[0, :, :, 0], sc.intrs[0], sc.extrs[0], levels=range(4), want_valid=True)
coords, targets = synth.make_tracks(clouds[0].xyz, w["N"], 12, seed=73, valid=clouds[0].xyz4[..., 3] > 0)
coords = (coords + 0.01 * torch.randn(coords.shape, device=dev, generator=torch.Generator(device=dev).manual_seed(74))).contiguous()


def timed(fn, reps=10):
    for _ in range(3):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


tot_f = tot_s = 0.0
for l, c in enumerate(clouds):
    out = torch.empty(12, w["N"], 64, device=dev)
    tf = timed(lambda: ops.knn_corr([(c, c.fvec, 16)], targets, coords, out=out))
    ts = timed(lambda: ops.knn(c, coords, 16, idx_dtype=torch.int32, want_dist=False))
    tot_f += tf; tot_s += ts
    print(f"level {l}: P={c.xyz.shape[1]:8d}  fused knn_corr {tf:.4f} ms   search only (mvt_knn) {ts:.4f} ms")
lv = [(c, c.fvec, 16) for c in clouds]
out = torch.empty(12, w["N"], 256, device=dev)
print(f"sum of levels: fused {tot_f:.4f} ms, search only {tot_s:.4f} ms;   one launch, 4 levels: {timed(lambda: ops.knn_corr(lv, targets, coords, out=out)):.4f} ms")
