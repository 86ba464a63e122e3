"""Diagnostics (GPU): warm start of the searches from the previous iteration's neighbours (idx_hint), bench-like motion
(independent N(0, 1 cm) perturbations of the tracks per iteration).    python tools/diag_warm.py [--workload panoptic]"""
import argparse, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mvtracker_b200 import ops, synth
from mvtracker_b200.hotpath import WindowClouds

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="panoptic")
ap.add_argument("--sigma", type=float, default=0.01)
a = ap.parse_args()
w = synth.WORKLOADS[a.workload]
dev = "cuda"
sc = synth.make_scene(w["V"], w["H"], w["W"], S=12, C=128, zero_frac=0.05, seed=72, device=dev)
win = WindowClouds.build(sc.fmaps[0], sc.depths[0, :, :, 0], sc.intrs[0], sc.extrs[0])
base, targets = synth.make_tracks(win.clouds[0].xyz, w["N"], 12, seed=73, valid=win.clouds[0].xyz4[..., 3] > 0)
g = torch.Generator(device=dev).manual_seed(74)
coords = [(base + a.sigma * torch.randn(base.shape, device=dev, generator=g)).contiguous() for _ in range(4)]
outs = {}
for warm in (False, True, False, True):
    win.warm_start, win._idx = warm, None
    res = [torch.empty(12, w["N"], 256, device=dev) for _ in range(4)]
    for _ in range(3):
        for i in range(4):
            win.corr(targets, coords[i], out=res[i])
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(5):
        for i in range(4):
            win.corr(targets, coords[i], out=res[i])
    e1.record()
    torch.cuda.synchronize()
    outs[warm] = [r.clone() for r in res]
    print(f"{a.workload} sigma={a.sigma} warm_start={warm}: {e0.elapsed_time(e1) / 20:.4f} ms per iteration")
print("identical results:", all(torch.equal(x, y) for x, y in zip(outs[False], outs[True])))
