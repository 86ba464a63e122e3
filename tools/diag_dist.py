"""Per-stage timing of the view-sharded path (run under torchrun): python -m torch.distributed.run ... tools/diag_dist.py [workload]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.distributed as dist
from mvtracker_b200 import ops, synth
from mvtracker_b200.dist import ShardedWindow, view_range
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
w = synth.WORKLOADS[sys.argv[1] if len(sys.argv) > 1 else "panoptic"]
V, N = w["V"], w["N"]
sc = synth.make_scene(V, w["H"], w["W"], S=12, C=128, zero_frac=0.05, seed=72, device=dev)
fm, dp, K, E = sc.fmaps[0], sc.depths[0, :, :, 0].contiguous(), sc.intrs[0].contiguous(), sc.extrs[0].contiguous()
c0 = ops.cloud_build(fm, dp, K, E, levels=[0], want_fvec=False, want_spacing=False)[0]
coords, targets = synth.make_tracks(c0.xyz, N, 12, seed=73, valid=c0.xyz4[..., 3] > 0)
del c0
v0, v1 = view_range(rank, world, V)
mode = sys.argv[2] if len(sys.argv) > 2 else "tracks"
full = ops.cloud_build(fm, dp, K, E, levels=range(4), want_spacing=False)
want = ops.knn_corr([(c, c.fvec, 16) for c in full], targets, coords)
del full
def build(reuse=None):
    return ShardedWindow.build(fm[v0:v1].contiguous(), dp[v0:v1].contiguous(), K[v0:v1].contiguous(), E[v0:v1].contiguous(), v0, V,
                               intrs_all=K, extrs_all=E, mode=mode, reuse=reuse)
win = build()
torch.cuda.synchronize(); dist.barrier()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5): win = build(win)
e1.record(); torch.cuda.synchronize()
print(f"rank {rank}: build ({mode}) {e0.elapsed_time(e1) / 5:.3f} ms", flush=True)
out = torch.empty(12, N, 256, device=dev)
for _ in range(3):
    win.corr(targets, coords, out=out)
torch.cuda.synchronize(); dist.barrier()
print(f"rank {rank}: sharded output equals the single-GPU fused kernel: {torch.equal(out, want)}", flush=True)
names = (["phaseA(keys+a2a+merge+gather)", "knn_keys", "all_to_all", "topk_merge", "all_gather", "corr_from_idx", "all_reduce"] if mode == "views"
         else ["search", "keys->idx", "all_gather", "corr_from_idx", "all_reduce"] if mode == "tracks"
         else ["search+idx stores", "barrier", "corr_from_idx+record stores", "barrier"])
acc = [0.0] * len(names)
R = 10
for _ in range(R):
    marks = []
    e0 = torch.cuda.Event(enable_timing=True); e0.record()
    win.corr(targets, coords, out=out, marks=marks)
    torch.cuda.synchronize()
    prev = e0
    for i, e in enumerate(marks):
        acc[i] += prev.elapsed_time(e); prev = e
print(f"rank {rank}: " + "  ".join(f"{n}={a / R:.3f}ms" for n, a in zip(names, acc)) + f"  total={sum(acc) / R:.3f}ms", flush=True)
# back-to-back calls: GPU time per call (events around the loop) and host time per call (how long Python takes to issue one)
import time
torch.cuda.synchronize(); dist.barrier()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(); t0 = time.perf_counter()
for _ in range(40):
    win.corr(targets, coords)
t1 = time.perf_counter(); e1.record(); torch.cuda.synchronize()
print(f"rank {rank}: back-to-back corr: {e0.elapsed_time(e1) / 40:.3f} ms per call on the GPU, {(t1 - t0) / 40 * 1e3:.3f} ms of host time to issue one", flush=True)
# raw NCCL timings at these sizes
for name, fn in (("all_reduce 50MB", lambda: dist.all_reduce(out)),):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): fn()
    e1.record(); torch.cuda.synchronize()
    if rank == 0: print(name, e0.elapsed_time(e1) / 10, "ms", flush=True)
dist.barrier(); dist.destroy_process_group()
