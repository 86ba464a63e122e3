"""Diagnostics (GPU): cloud_build timing variants for a workload.   python tools/diag_cloud.py --workload droid"""
import argparse, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mvtracker_b200 import ops, synth

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="droid")
ap.add_argument("--S", type=int, default=12)
ap.add_argument("--reps", type=int, default=20)
a = ap.parse_args()
w = synth.WORKLOADS[a.workload]
sc = synth.make_scene(w["V"], w["H"], w["W"], S=a.S, C=128, zero_frac=0.05, seed=72, device="cuda")
fm, dp, K, E = sc.fmaps[0], sc.depths[0, :, :, 0].contiguous(), sc.intrs[0], sc.extrs[0]
cam = ops.camera_prep(K, E)


def timeit(tag, **kw):
    for _ in range(3):
        c = ops.cloud_build(fm, dp, K, E, cam=cam, **kw)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.reps):
        c = ops.cloud_build(fm, dp, K, E, cam=cam, **kw)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / a.reps
    nbytes = fm.numel() * fm.element_size() + dp.numel() * 4
    for x in c:
        for t in (x.xyz, x.fvec, x.xyz4, x.valid):
            if t is not None:
                nbytes += t.numel() * t.element_size()
    print(f"{tag:40s} {ms:7.3f} ms  {nbytes / ms / 1e6:8.1f} GB/s (all bytes moved)")


timeit("levels 0-3, xyz4+spacing", levels=range(4))
timeit("levels 0-3, no xyz4", levels=range(4), want_xyz4=False)
timeit("levels 0-3, no xyz4, no spacing", levels=range(4), want_xyz4=False, want_spacing=False)
timeit("levels 0-3, no fvec", levels=range(4), want_fvec=False)
timeit("level 0 only", levels=[0])
timeit("level 0 only, no xyz4/spacing", levels=[0], want_xyz4=False, want_spacing=False)
timeit("levels 0-3 + valid", levels=range(4), want_valid=True)
fmh = fm.to(torch.bfloat16)
fm = fmh
timeit("bf16 levels 0-3", levels=range(4))
