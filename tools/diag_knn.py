"""Diagnostics (GPU): grid parameters and search statistics of the kNN kernel for a workload / cell scale sweep.
    python tools/diag_knn.py --workload droid --scales 0.5,1,2,4
"""
import argparse, ctypes, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from mvtracker_b200 import _lib, ops, synth

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="droid")
ap.add_argument("--scales", default="1")
ap.add_argument("--S", type=int, default=12)
ap.add_argument("--zero-frac", type=float, default=0.05)
ap.add_argument("--jitter", type=float, default=0.02)
ap.add_argument("--static", action="store_true", help="tracks fixed at their frame-0 position (drift off moving surfaces)")
a = ap.parse_args()
w = synth.WORKLOADS[a.workload]
dev = "cuda"
sc = synth.make_scene(w["V"], w["H"], w["W"], S=a.S, C=128, zero_frac=a.zero_frac, seed=72, device=dev)
clouds = ops.cloud_build(sc.fmaps[0], sc.depths[0, :, :, 0], sc.intrs[0], sc.extrs[0], levels=range(4), want_valid=True)
coords, targets = synth.make_tracks(clouds[0].xyz if not a.static else clouds[0].xyz[0], w["N"], a.S, seed=73, jitter=a.jitter,
                                    valid=clouds[0].valid if not a.static else clouds[0].valid[0])
L = _lib.lib()
stats = torch.zeros(16, dtype=torch.int64, device=dev)
names = ["queries", "ring1", "ring2", "ring3", "fallback", "cand", "ins", "lookups", "small"]
def run(tag, levels_of):
    for lvl, c in enumerate(clouds):
        L.mvt_debug_set_stats(ctypes.c_void_p(stats.data_ptr()))
        stats.zero_()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ops.knn_corr([(levels_of(lvl), c.fvec, 16)], targets, coords)
        e1.record()
        torch.cuda.synchronize()
        L.mvt_debug_set_stats(None)
        st = stats.cpu().tolist()
        q = max(st[0], 1)
        print(f"{tag} lvl {lvl} N={c.xyz.shape[1]:7d} | ms={e0.elapsed_time(e1):.3f} pass1={st[1] / q:.3f} pass2={st[2] / q:.3f} "
              f"pass3={st[3] / q:.3f} pass4+={st[4] / q:.4f} cand/q={st[5] / q:.0f} ins/q={st[6] / q:.1f} "
              f"hist(<2,<16,<128,<1k,<8k,>=8k)={st[10:16]} tiles tested/q={st[7] / q:.2f} kept/q={st[8] / q:.2f}")
        worst_c, worst_q = st[9] >> 24, st[9] & 0xFFFFFF
        b, m_ = divmod(worst_q, coords.shape[1])
        qq = coords[b, m_]
        cam = c.cam.view(c.V, -1, 48)[:, b]
        R, t = cam[:, 24:33].view(-1, 3, 3), cam[:, 33:36]
        qc = (R @ qq) + t
        d, i = ops.knn(ops.Search.from_cloud(c, frame=b), qq[None, None], 16)
        print(f"    worst query: {worst_c} candidates, frame {b} track {m_} q={[round(x, 3) for x in qq.tolist()]} "
              f"q_cam per view={[[round(x, 2) for x in r] for r in qc.tolist()]} d16={d[0, 0, -1].item():.4f} d1={d[0, 0, 0].item():.4f}")


run("raster", lambda l: clouds[l])
for sc_ in [float(x) for x in a.scales.split(",")]:
    grids = ops.build_grids([(c.xyz, None, c.spacing) for c in clouds], k_hint=16, cell_scale=sc_)
    for lvl, (g, c) in enumerate(zip(grids, clouds)):
        info = g.info().cpu()
        i0 = info[0]
        L.mvt_debug_set_stats(ctypes.c_void_p(stats.data_ptr()))
        stats.zero_()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ops.knn_corr([(g, c.fvec, 16)], targets, coords)
        e1.record()
        torch.cuda.synchronize()
        L.mvt_debug_set_stats(None)
        st = stats.cpu().tolist()
        q = max(st[0], 1)
        print(f"scale {sc_:4.2f} lvl {lvl} N={g.N:7d} h={i0[3]:.4f} dims={int(i0[4])}x{int(i0[5])}x{int(i0[6])} cells={int(i0[8])} "
              f"pts/cell={i0[7] / max(i0[8], 1):.1f} | ms={e0.elapsed_time(e1):.3f} ring1={st[1] / q:.3f} ring2={st[2] / q:.3f} "
              f"ring3={st[3] / q:.3f} fallback={st[4] / q:.4f} small={st[8] / q:.3f} cand/q={st[5] / q:.0f} ins/q={st[6] / q:.1f} "
              f"lookups/q={st[7] / q:.0f}")

# fused launch, stats off, back-to-back (true kernel time) and host cost of one call
import time
for tag, lv in (("raster fused x4", [(c, c.fvec, 16) for c in clouds]), ("raster lvl0 only", [(clouds[0], clouds[0].fvec, 16)])):
    out = ops.knn_corr(lv, targets, coords)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    for _ in range(20):
        ops.knn_corr(lv, targets, coords, out=out)
    e1.record()
    t_host = (time.perf_counter() - t0) / 20
    torch.cuda.synchronize()
    print(f"{tag}: {e0.elapsed_time(e1) / 20:.3f} ms per launch (host {t_host * 1e3:.3f} ms per call)")
