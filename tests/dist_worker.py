"""Worker of tests/test_gpu_torchrun.py (run under torchrun, one rank per GPU): every transport of mvtracker_b200.dist must
reproduce the single-GPU fused kernel bit for bit on real ranks -- in particular "fused" (mvt_shard_iter: peer stores into
the owners' inboxes, multicast result stores, in-kernel epoch exchange) and "p2p", which the emulated-rank tests cannot run."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from mvtracker_b200 import ops, synth
from mvtracker_b200.dist import ShardedWindow, view_range

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
V, Hs, Ws, S, N = 4, 48, 64, 12, 517
ok = True
for dtype in (torch.float32, torch.bfloat16):
    sc = synth.make_scene(V, Hs * 4, Ws * 4, S=S, C=128, zero_frac=0.1, seed=72, device=dev, dtype=dtype)
    fm, dp, K, E = sc.fmaps[0], sc.depths[0, :, :, 0].contiguous(), sc.intrs[0].contiguous(), sc.extrs[0].contiguous()
    full = ops.cloud_build(fm, dp, K, E, levels=range(4), want_spacing=False)
    v0, v1 = view_range(rank, world, V)
    for mode in ("fused", "p2p", "tracks"):
        win = None
        for it in range(3):                       # three windows / iterations through the same buffers
            coords, targets = synth.make_tracks(full[0].xyz, N, S, seed=73 + it, dtype=dtype, valid=full[0].xyz4[..., 3] > 0)
            coords[:, :11] += 1.5                 # free-space queries
            want = ops.knn_corr([(c, c.fvec, 16) for c in full], targets, coords, defer=False)
            win = ShardedWindow.build(fm[v0:v1].contiguous(), dp[v0:v1].contiguous(), K[v0:v1].contiguous(), E[v0:v1].contiguous(),
                                      v0, V, intrs_all=K, extrs_all=E, mode=mode, reuse=win)
            for rep in range(2):
                got = win.corr(targets, coords)
                torch.cuda.synchronize()
                same = torch.equal(got, want)
                err = win._bufs["shard_state"].error() if mode == "fused" and "shard_state" in win._bufs else False
                if not same or err or win.mode != mode:
                    ok = False
                    print(f"rank {rank}: MISMATCH mode={mode} (ran as {win.mode}) dtype={dtype} window {it} rep {rep}: "
                          f"{(got != want).any(-1).sum().item()} rows differ, timeout flag {err}", flush=True)
        dist.barrier()
# mode "frames": frames sharded round robin, every rank's fused kernel stores into every rank's block (multimem.st / peer stores)
from mvtracker_b200.dist import FrameShardedSequence
T, hop = 24, 6
for dtype in (torch.float32, torch.bfloat16):
    sc = synth.make_scene(V, Hs * 4, Ws * 4, S=T, C=128, zero_frac=0.1, seed=75, device=dev, dtype=dtype)
    fm, dp, K, E = sc.fmaps[0], sc.depths[0, :, :, 0].contiguous(), sc.intrs[0].contiguous(), sc.extrs[0].contiguous()
    seq = FrameShardedSequence(V, Hs, Ws, 128, dtype, dev, S=S, hop=hop)
    for wdw in range(3):
        win = slice(hop * wdw, hop * wdw + S)
        full = ops.cloud_build(fm[:, win].contiguous(), dp[:, win].contiguous(), K[:, win].contiguous(), E[:, win].contiguous(),
                               levels=range(4), want_spacing=False)
        coords, targets = synth.make_tracks(full[0].xyz, N, S, seed=80 + wdw, dtype=dtype, valid=full[0].xyz4[..., 3] > 0)
        coords[:, :11] += 1.5
        want = ops.knn_corr([(c, c.fvec, 16) for c in full], targets, coords, defer=False)
        if wdw == 0:
            ts = seq.new_frames()
            seq.push(fm[:, ts].contiguous() if ts else None, dp[:, ts].contiguous() if ts else None, K[:, win], E[:, win])
        else:
            seq.push()                                  # the window that was prefetched during the previous window's iterations
        for rep in range(3):
            got = seq.corr(targets, coords)
            if rep == 0 and wdw < 2:                    # build the next window's frames on the side stream, under the iterations
                ts = seq.new_frames()
                nxt = slice(hop * (wdw + 1), hop * (wdw + 1) + S)
                held = (fm[:, ts].contiguous() if ts else None, dp[:, ts].contiguous() if ts else None, K[:, nxt], E[:, nxt])
                seq.prefetch(*held)
            torch.cuda.synchronize()
            if not torch.equal(got, want) or seq.transport != "multicast":
                ok = False
                print(f"rank {rank}: MISMATCH mode=frames ({seq.transport}) dtype={dtype} window {wdw} rep {rep}: "
                      f"{(got != want).any(-1).sum().item()} rows differ", flush=True)
    dist.barrier()
t = torch.tensor([0 if ok else 1], device=dev)
dist.all_reduce(t)
if rank == 0:
    print("DIST_WORKER_OK" if t.item() == 0 else "DIST_WORKER_FAILED", flush=True)
dist.barrier()
dist.destroy_process_group()
sys.exit(0 if t.item() == 0 else 1)
