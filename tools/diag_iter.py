"""Per-iteration time of the sharded correlation on real ranks (torchrun, one rank per GPU), per transport:
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 tools/diag_iter.py [workload] [modes]
Timed on the device (CUDA events around `reps` back-to-back iterations), max over ranks; the output of every mode is compared
with the single-GPU fused kernel.  Knobs of mvt_shard_iter come from the environment (MVT_SHARD_EARLY_PCT, MVT_SHARD_DEFER,
MVT_SHARD_NO_LOCAL)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from mvtracker_b200 import ops, synth
from mvtracker_b200.dist import FrameShardedSequence, ShardedWindow, view_range

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
name = sys.argv[1] if len(sys.argv) > 1 else "panoptic"
modes = sys.argv[2].split(",") if len(sys.argv) > 2 else ["frames", "fused", "p2p"]
w = synth.WORKLOADS[name]
V, N, S = w["V"], w["N"], 12
sc = synth.make_scene(V, w["H"], w["W"], S=S, C=128, zero_frac=0.05, seed=72, device=dev)
fm, dp, K, E = sc.fmaps[0], sc.depths[0, :, :, 0].contiguous(), sc.intrs[0].contiguous(), sc.extrs[0].contiguous()
full = ops.cloud_build(fm, dp, K, E, levels=range(4), want_spacing=False)
coords, targets = synth.make_tracks(full[0].xyz, N, S, seed=73, valid=full[0].xyz4[..., 3] > 0)
want = ops.knn_corr([(c, c.fvec, 16) for c in full], targets, coords, defer=False)
v0, v1 = view_range(rank, world, V)
knobs = " ".join(f"{k[10:].lower()}={os.environ[k]}" for k in ("MVT_SHARD_EARLY_PCT", "MVT_SHARD_DEFER", "MVT_SHARD_NO_LOCAL") if k in os.environ)
for mode in modes:
    if mode == "frames":
        win = FrameShardedSequence(V, w["H"] // 4, w["W"] // 4, 128, torch.float32, dev, S=S)
        ts = win.new_frames()
        win.push(fm[:, ts].contiguous(), dp[:, ts].contiguous(), K, E)
    else:
        win = ShardedWindow.build(fm[v0:v1].contiguous(), dp[v0:v1].contiguous(), K[v0:v1].contiguous(), E[v0:v1].contiguous(),
                                  v0, V, intrs_all=K, extrs_all=E, mode=mode)
    for _ in range(3):
        got = win.corr(targets, coords)
    torch.cuda.synchronize()
    same = torch.equal(got, want)
    dist.barrier()
    reps = 20
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        win.corr(targets, coords)
    b.record()
    torch.cuda.synchronize()
    t = torch.tensor([a.elapsed_time(b) / reps, 0.0 if same else 1.0], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        print(f"N={world} {name} {win.mode:6s} {knobs}: {t[0].item():.3f} ms per iteration, equal {t[1].item() == 0}", flush=True)
    del win
    dist.barrier()
dist.destroy_process_group()
