"""Search statistics of ONE rank's local search in the view-sharded path (single GPU): python tools/diag_local.py [world]"""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mvtracker_b200 import _lib, ops, synth
from mvtracker_b200.dist import view_range
dev = "cuda"
world = int(sys.argv[1]) if len(sys.argv) > 1 else 2
w = synth.WORKLOADS["panoptic"]
V, N, S = w["V"], w["N"], 4
sc = synth.make_scene(V, w["H"], w["W"], S=S, C=128, zero_frac=0.05, seed=72, device=dev)
fm, dp, K, E = sc.fmaps[0], sc.depths[0, :, :, 0].contiguous(), sc.intrs[0].contiguous(), sc.extrs[0].contiguous()
full = ops.cloud_build(fm, dp, K, E, levels=range(4), want_spacing=False)
coords, targets = synth.make_tracks(full[0].xyz, N, S, seed=73, valid=full[0].xyz4[..., 3] > 0)
L = _lib.lib()
stats = torch.zeros(16, dtype=torch.int64, device=dev)
def timed(fn, tag):
    L.mvt_debug_set_stats(ctypes.c_void_p(stats.data_ptr())); stats.zero_(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); r = fn(); e1.record(); torch.cuda.synchronize()
    L.mvt_debug_set_stats(None)
    st = stats.cpu().tolist(); q = max(st[0], 1)
    print(f"{tag}: {e0.elapsed_time(e1):.3f} ms  passes(1,2,3,4+)={[round(x / q, 3) for x in st[1:5]]} cand/q={st[5] / q:.0f} ins/q={st[6] / q:.1f} "
          f"hist(<2,<16,<128,<1k,<8k,>=8k)={st[10:16]} worst={st[9] >> 24} cand (query {st[9] & 0xFFFFFF})", flush=True)
    return r
for l in range(4):
    timed(lambda: ops.knn_keys([full[l]], [0], coords, 16), f"all {V} views, level {l}")
for r in range(world):
    v0, v1 = view_range(r, world, V)
    cl = ops.cloud_build(fm[v0:v1].contiguous(), dp[v0:v1].contiguous(), K[v0:v1].contiguous(), E[v0:v1].contiguous(), levels=range(4), want_spacing=False)
    for l in range(4):
        timed(lambda: ops.knn_keys([cl[l]], [v0 * cl[l].H * cl[l].W], coords, 16), f"rank {r} views {v0}..{v1 - 1}, level {l}")
print("--- two phases on rank 0's views (caps from the merged phase-A lists of all ranks)")
locs = []
for r in range(world):
    v0, v1 = view_range(r, world, V)
    locs.append((v0, ops.cloud_build(fm[v0:v1].contiguous(), dp[v0:v1].contiguous(), K[v0:v1].contiguous(), E[v0:v1].contiguous(), levels=range(4), want_spacing=False)))
for budget in [int(b) for b in (sys.argv[2] if len(sys.argv) > 2 else "120,360,720,2000").split(",")]:
    for l in range(4):
        fl = [torch.empty(1, 1, S, N, dtype=torch.uint8, device=dev) for _ in locs]
        kb = [torch.empty(1, 1, S, N, 16, dtype=torch.int64, device=dev) for _ in locs]
        for i, (v0, cl) in enumerate(locs):
            f = lambda: ops.knn_keys([cl[l]], [v0 * cl[l].H * cl[l].W], coords, 16, budget=budget, exact_out=fl[i], out=kb[i])
            if i == 0: timed(f, f"budget {budget} level {l} phase A rank 0")
            else: f()
        merged = torch.empty(S * N, 16, dtype=torch.int64, device=dev)
        ops.topk_merge(torch.stack([b[0, 0] for b in kb]).view(world, S * N, 16).contiguous(), keys_out=merged)
        cap = merged[:, 15].contiguous()
        v0, cl = locs[0]
        timed(lambda: ops.knn_keys([cl[l]], [0], coords, 16, caps=cap, skip=fl[0], out=kb[0]),
              f"budget {budget} level {l} phase B rank 0 (converged in A: {fl[0].float().mean():.3f})")
