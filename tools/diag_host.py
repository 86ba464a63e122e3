"""Host (Python + launch) time against GPU time of one step of mode "frames", emulating rank 0 of `world` ranks on ONE GPU
(stores go to a local block, no barrier): is a step at N = 8 bound by the host?   python tools/diag_host.py [world] [workload]"""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mvtracker_b200 import ops, synth
from mvtracker_b200.dist import FrameShardedSequence
world = int(sys.argv[1]) if len(sys.argv) > 1 else 8
w = synth.WORKLOADS[sys.argv[2] if len(sys.argv) > 2 else "panoptic"]
dev = torch.device("cuda", 0)
V, N, S, HOP, POOL = w["V"], w["N"], 12, 6, 24
sc = synth.make_scene(V, w["H"], w["W"], S=POOL // world, C=128, zero_frac=0.05, seed=72, device=dev)     # rank 0's pool frames
fm, dp = sc.fmaps[0], sc.depths[0, :, :, 0].contiguous()
K = sc.intrs[0][:, :1].expand(V, S, 3, 3).contiguous(); E = sc.extrs[0][:, :1].expand(V, S, 3, 4).contiguous()
c0 = ops.cloud_build(fm[:, :1].contiguous(), dp[:, :1].contiguous(), K[:, :1].contiguous(), E[:, :1].contiguous(), levels=[0], want_fvec=False)[0]
coords1, targets1 = synth.make_tracks(c0.xyz, N, 1, seed=73, valid=c0.xyz4[..., 3] > 0)
coords, targets = coords1.expand(S, N, 3).contiguous(), targets1.expand(S, N, 128).contiguous()
out = torch.empty(S, N, 256, device=dev)
seq = FrameShardedSequence(V, w["H"] // 4, w["W"] // 4, 128, torch.float32, dev, S=S, rank=0, world=world, shared_out=out)
t_push = t_corr = 0.0
n_steps, iters = 24, 4
def step(timed):
    global t_push, t_corr
    ts = seq.new_frames()
    a = time.perf_counter()
    li = [((t % POOL) - 0) // world for t in ts]
    seq.push(fm[:, li[0]:li[0] + len(li)] if ts else None, dp[:, li[0]:li[0] + len(li)] if ts else None, K, E)
    b = time.perf_counter()
    for _ in range(iters):
        seq.corr(targets, coords)
    c = time.perf_counter()
    if timed:
        t_push += b - a; t_corr += c - b
for _ in range(4): step(False)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
h0 = time.perf_counter(); e0.record()
for _ in range(n_steps): step(True)
e1.record(); h1 = time.perf_counter()
torch.cuda.synchronize()
print(f"world {world} rank 0 emulated: host enqueue {1e3 * (h1 - h0) / n_steps:.3f} ms/step (push {1e3 * t_push / n_steps:.3f}, "
      f"{iters} x corr {1e3 * t_corr / n_steps / iters:.3f} each) | GPU {e0.elapsed_time(e1) / n_steps:.3f} ms/step", flush=True)
