"""knn_keys (search only) vs the fused knn_corr kernel on one GPU, same queries: python tools/diag_keys.py [workload] [frac]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mvtracker_b200 import ops, synth
dev = "cuda"
w = synth.WORKLOADS[sys.argv[1] if len(sys.argv) > 1 else "panoptic"]
V, N, S = w["V"], w["N"], 12
sc = synth.make_scene(V, w["H"], w["W"], S=S, C=128, zero_frac=0.05, seed=72, device=dev)
full = ops.cloud_build(sc.fmaps[0], sc.depths[0, :, :, 0].contiguous(), sc.intrs[0].contiguous(), sc.extrs[0].contiguous(), levels=range(4), want_spacing=False)
coords, targets = synth.make_tracks(full[0].xyz, N, S, seed=73, valid=full[0].xyz4[..., 3] > 0)
def t(fn, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
lv = [(c, c.fvec, 16) for c in full]
if os.environ.get("MVT_GROUP_WIDTH"):
    from mvtracker_b200 import _lib
    _lib.lib().mvt_debug_set_group_width(int(os.environ["MVT_GROUP_WIDTH"]))
out = torch.empty(S, N, 256, device=dev)
print(f"fused knn_corr, {S * N * 4} queries: {t(lambda: ops.knn_corr(lv, targets, coords, out=out)):.3f} ms")
for frac in (1, 2, 4, 8):
    m = N // frac
    c = coords[:, :m].contiguous()
    keys = torch.empty(1, 4, S, m, 16, dtype=torch.int64, device=dev)
    print(f"knn_keys, {S * m * 4} queries (1/{frac}): {t(lambda: ops.knn_keys(full, [0] * 4, c, 16, out=keys)):.3f} ms")
# corr_from_idx with the ownership of one rank out of n (views 0 .. V/n - 1), all queries
_, fidx = ops.knn_corr(lv, targets, coords, return_idx=True)
idx_all = fidx.unsqueeze(0).contiguous()            # [1 shard, L, S, M, k]
for n in (1, 2, 4, 8):
    owned = [(c.xyz4[:, :(V // n) * c.H * c.W].contiguous(), c.fvec[:, :(V // n) * c.H * c.W].contiguous(), 0, (V // n) * c.H * c.W) for c in full]
    res = torch.empty(S, N, 256, device=dev)
    print(f"corr_from_idx, owner of {V // n}/{V} views: {t(lambda: ops.corr_from_idx(owned, idx_all, targets, coords, 16, out=res)):.3f} ms")
