#!/usr/bin/env python
"""bench.py -- kNN-corr track-point queries/s of the MVTracker hot path on B200 (BASELINE.json metric).

One STEP = the hot path of one sliding window (S=12 frames) on synthetic inputs already resident in HBM:
    camera_prep + cloud_build (4 pyramid levels, one pass)  [-> grid_build (4 levels) with --search grid]  ->
    iters x knn_corr (all levels / frames / tracks fused in one launch; raster search needs no index)
queries per step = iters * levels * S * N_tracks  (one query = one (track, frame, level, iteration) tuple).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload droid|panoptic|dexycb|kubric] [--impl reference]
N > 1 is launched by torchrun (one rank per GPU); see mvtracker_b200/dist.py for the sharding.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

S_WINDOW, C_FEAT, K_NN, LEVELS, STRIDE = 12, 128, 16, 4, 4
# The workload every N runs by default: the 8-view / ~1M-point window north_star quotes its 1-GPU roofline and 8-GPU scaling
# targets on (BASELINE.json configs[2]); it fits one GPU (~17 GB), so the 1/2/4/8-GPU series is one fixed job (strong scaling).
DEFAULT_WORKLOAD = "panoptic"


def ncu_traffic(workload, dtype, kernel, n_gpus):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of `kernel`, from the ncu --set full capture that
    tools/ncu_summary.py recorded in profiles/traffic.json ("workload/dtype/kernel/nN" -> {bytes, capture}); None when that
    kernel has no capture at this N (a number measured on another kernel or another N is not this kernel's traffic)."""
    try:
        tab = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        return tab.get(f"{workload}/{dtype}/{kernel}/n{n_gpus}", {}).get("bytes")
    except Exception:
        return None


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=None)
    ap.add_argument("--iters", type=int, default=4)
    ap.add_argument("--dtype", default="f32", choices=["f32", "bf16"])
    ap.add_argument("--zero-frac", type=float, default=0.05)
    ap.add_argument("--cell-scale", type=float, default=0.0)
    ap.add_argument("--search", default="raster", choices=["raster", "grid"],
                    help="raster: search the depth rasters directly (no index build); grid: hashed grid")
    ap.add_argument("--dist-mode", default="frames", choices=["frames", "fused", "p2p", "tracks", "views"],
                    help="N > 1: frames (default) = the window's frames are sharded round robin, each rank runs the fused kernel on its "
                         "frames x all tracks and stores the records into every rank's block; fused / p2p = views sharded for the build, "
                         "tracks for the search (one persistent kernel / four stages per iteration); tracks / views = NCCL transports")
    ap.add_argument("--scaling", default="strong", choices=["strong", "weak"],
                    help="N > 1: strong = ONE window sharded over the ranks (views, then tracks; the default, what the path's "
                         "multi-GPU design is about); weak = every rank tracks its own sequence's window with the single-GPU path "
                         "(independent replicas, no exchange)")
    ap.add_argument("--step", default="advance", choices=["advance", "window"],
                    help="advance (default): a step is ONE WINDOW ADVANCE of a sequence -- the 6 new frames are built into the "
                         "per-frame cloud cache (the other 6 are shared with the previous window) + iters x knn_corr over the 12-frame "
                         "window; window: every step rebuilds all 12 frames, as the reference does (mvtracker.py:297-306)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--warm-start", action="store_true",
                    help="N = 1: iterations 2..4 of a window start their searches from the previous iteration's neighbours "
                         "(exact; off by default so that every timed iteration does the full cold search)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-prefetch", action="store_true",
                    help="N > 1, --dist-mode frames: build a window's new frames at the start of its step instead of on a side stream "
                         "during the previous window's iterations")
    ap.add_argument("--no-verify", action="store_true", help="skip the sampled check of the last step's output against the oracle")
    ap.add_argument("--no-distributions", action="store_true",
                    help="skip the extra timed track distributions (static init, 40 %% invalid depth); N = 1 only")
    ap.add_argument("--no-model", action="store_true", help="skip the MVTracker.forward frames/s legs (N = 1 only)")
    ap.add_argument("--model-workloads", default="kubric,droid", help="configs of the MVTracker.forward legs")
    ap.add_argument("--no-adjacent", action="store_true", help="skip the timings of the adjacent kernels (rows f1 / f2)")
    ap.add_argument("--cpu-budget", type=float, default=12.0, help="seconds of CPU work for the cpu_baseline sample")
    return ap.parse_args()


def workload_cfg(name):
    """A named workload (synth.WORKLOADS), or a point of the scaling sweep: "sweep:V=8,P0=1M,N=16k" (V views, P0 fused
    level-0 points per frame, N tracks; the raster is 4:5, as the panoptic one)."""
    from mvtracker_b200 import synth
    if name.startswith("sweep:"):
        kv = dict(item.split("=") for item in name[6:].split(","))
        num = lambda t: int(float(t[:-1]) * {"k": 1024, "M": 1024 * 1024}[t[-1]]) if t[-1] in "kM" else int(t)
        V, P0, N = int(kv["V"]), num(kv["P0"]), num(kv["N"])
        hs = max(8, int(round((P0 / V / 1.25) ** 0.5 / 8)) * 8)          # level-0 raster rows (multiple of 8: 4 levels)
        ws = max(8, int(round(P0 / V / hs / 8)) * 8)
        w = {"V": V, "H": hs * STRIDE, "W": ws * STRIDE, "N": N, "T": 24}
    else:
        w = dict(synth.WORKLOADS[name])
    w["name"] = name
    w["Hs"], w["Ws"] = w["H"] // STRIDE, w["W"] // STRIDE
    return w


# ------------------------------------------------------------------------------------------------------------------
# clocks: sample nvidia-smi DURING the timed region
# ------------------------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.samples, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.samples.append((time.time(), line.strip()))

    def window(self, t0, t1):
        sm, mx, reasons = [], 0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ts, line in self.samples:
            if ts < t0 - 0.05 or ts > t1 + 0.15:
                continue
            parts = [p.strip() for p in line.split(",")]
            try:
                sm.append(float(parts[0])); mx = max(mx, float(parts[1]))
            except Exception:
                continue
            for n, v in zip(names, parts[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons),
                "samples": len(sm)}

    def stop(self):
        if self.proc:
            self.proc.terminate()


# ------------------------------------------------------------------------------------------------------------------
# CPU baseline: the oracle port of the reference path, all host threads, bounded sample of the same workload
# ------------------------------------------------------------------------------------------------------------------
_scene_cache = {}


def cpu_sample(w, iters, budget_s, zero_frac, seed=72):
    """Times, on ONE frame of the workload: unproject x4 levels (re-pooling from level 0 per level, as
    model_utils.py:439-446 does) and corr_sample x4 levels on a track subsample sized to the budget; cost is exactly
    linear in frames and tracks, so the whole-step CPU time is S*t_build + iters*S*N*sum_l t_query_l.  -> dict"""
    import numpy as np
    import torch
    from mvtracker_b200 import synth
    from oracle import oracle

    # all the host threads this process may use -- set explicitly: torchrun exports OMP_NUM_THREADS=1 to its workers,
    # which would time the CPU arm on one core
    oracle.set_num_threads(len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1))
    cores = oracle.num_threads()
    ck = (w["name"], zero_frac, seed)
    if _scene_cache.get("key") != ck:       # inputs are generated once; only the oracle calls are timed
        sc = synth.make_scene(w["V"], w["H"], w["W"], S=1, C=C_FEAT, stride=STRIDE, zero_frac=zero_frac, seed=seed)
        _scene_cache.update(key=ck, arrays=(sc.fmaps[0].numpy(), sc.depths[0, :, :, 0].numpy(), sc.intrs[0].numpy(),
                                            sc.extrs[0].numpy()))
    fm, dp, K, E = _scene_cache["arrays"]
    t0 = time.perf_counter()
    clouds = [oracle.unproject(fm, dp, K, E, stride=STRIDE, level=l) for l in range(LEVELS)]
    t_build = time.perf_counter() - t0
    coords, targets = synth.make_tracks(torch.from_numpy(clouds[0][0][0]), w["N"], 1, C=C_FEAT)
    coords, targets = coords.numpy(), targets.numpy()

    def run(m):
        t = time.perf_counter()
        for xyz, fvec in clouds:
            oracle.corr_sample(K_NN, xyz, fvec, targets[:, :m], coords[:, :m])
        return time.perf_counter() - t

    m = min(w["N"], 16)
    t_small = run(m)
    per_track = t_small / m
    m_big = int(max(m, min(w["N"], (budget_s - t_build - t_small) / max(per_track, 1e-9))))
    reps = 1
    if m_big >= w["N"]:                       # whole track set fits the budget: repeat it to fill the sample
        reps = int(max(1, min(64, (budget_s - t_build - t_small) / max(per_track * m_big, 1e-9))))
    t_corr = sum(run(m_big) for _ in range(reps)) / reps
    t_q = t_corr / m_big                      # seconds per track per frame, all 4 levels
    step_cpu = S_WINDOW * t_build + iters * S_WINDOW * w["N"] * t_q
    queries = iters * LEVELS * S_WINDOW * w["N"]
    return {"value": queries / step_cpu, "unit": "queries/s", "cores": cores, "kind": "port",
            "sample": (f"oracle (C/OpenMP port of the reference path) on 1 of {S_WINDOW} frames: unproject x{LEVELS} levels "
                       f"({t_build:.2f}s) + corr_sample x{LEVELS} levels on {m_big} of {w['N']} tracks ({t_corr:.2f}s, mean of {reps}); "
                       f"scaled linearly to {S_WINDOW} frames x {w['N']} tracks x {iters} iters"),
            "step_seconds_extrapolated": step_cpu, "sample_seconds": t_build + t_small + t_corr * reps}


def ref_sample(w, iters, budget_s, zero_frac, seed=72):
    """The reference's OWN CPU path (oracle/ref_import.py: /root/reference, or the unmodified copy staged under oracle/_ref):
    init_pointcloud_from_rgbd x4 levels (model_utils.py:420-482) and PointcloudCorrBlock.corr_sample -> _knn_torch
    (mvtracker.py:75-79, 800-846) x4 levels x `iters`, on ONE frame of the workload and a track subsample sized to the budget
    (_knn_torch materialises [S, M, P] fp32 -- 201 GB at the panoptic level 0 -- so it is run per frame and per
    track chunk; its cost is exactly linear in both).  Everything reported under sample_* is measured; `value` scales the
    measured per-frame build and per-track correlation times to the window (S frames, N tracks)."""
    import numpy as np
    import torch
    from mvtracker_b200 import synth
    from oracle import ref_import
    ref = ref_import.load()
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    torch.set_num_threads(cores)
    ck = (w["name"], zero_frac, seed, "ref")
    if _scene_cache.get("key") != ck:
        sc = synth.make_scene(w["V"], w["H"], w["W"], S=1, C=C_FEAT, stride=STRIDE, zero_frac=zero_frac, seed=seed)
        _scene_cache.clear()
        _scene_cache.update(key=ck, scene=sc)
    sc = _scene_cache["scene"]
    with torch.no_grad():
        t0 = time.perf_counter()
        blocks = []
        for lvl in range(LEVELS):
            xyz, fvec = ref.init_pointcloud_from_rgbd(fmaps=sc.fmaps, depths=sc.depths, intrs=sc.intrs, extrs=sc.extrs,
                                                      stride=STRIDE, level=lvl)
            blocks.append(ref.PointcloudCorrBlock(k=K_NN, groups=1, xyz=xyz, fvec=fvec, corr_add_neighbor_offset=True,
                                                  corr_add_neighbor_xyz=False))
        t_build = time.perf_counter() - t0
        coords, targets = synth.make_tracks(blocks[0].xyz[0], w["N"], 1, C=C_FEAT)
        P0 = blocks[0].xyz.shape[1]
        chunk = int(max(16, min(w["N"], (2 << 30) // (4 * P0))))      # <= 2 GB distance matrix per call

        def run(m):
            t = time.perf_counter()
            for a in range(0, m, chunk):
                b = min(m, a + chunk)
                for blk in blocks:
                    blk.corr_sample(targets[:, a:b], coords[:, a:b])
            return time.perf_counter() - t

        m = min(w["N"], 16)
        t_small = run(m)
        per_track = t_small / m
        m_big = int(max(m, min(w["N"], (budget_s - t_build - t_small) / max(per_track * iters, 1e-9))))
        t_corr = sum(run(m_big) for _ in range(iters))         # `iters` refinement iterations on the same clouds
    t_q = t_corr / (iters * m_big)                              # seconds per track per frame per iteration, all 4 levels
    step_cpu = S_WINDOW * t_build + iters * S_WINDOW * w["N"] * t_q
    queries = iters * LEVELS * S_WINDOW * w["N"]
    sample_s = t_build + t_corr
    return {"value": queries / step_cpu, "unit": "queries/s", "cores": cores, "kind": "reference",
            "source": f"{ref.kind}: {'oracle/_ref (unmodified files staged by oracle/make_ref.py)' if ref.kind == 'staged' else ref.root}",
            "sample": (f"the reference's own init_pointcloud_from_rgbd x{LEVELS} levels ({t_build:.2f}s) + PointcloudCorrBlock.corr_sample "
                       f"(_knn_torch) x{LEVELS} levels x{iters} iterations on {m_big} of {w['N']} tracks ({t_corr:.2f}s), 1 of {S_WINDOW} frames, "
                       f"{cores} torch threads; value = window queries / (S*t_build + S*N/m*t_corr)"),
            "sample_queries": iters * LEVELS * m_big, "sample_seconds": sample_s,
            "sample_value": iters * LEVELS * m_big / sample_s, "step_seconds_extrapolated": step_cpu}


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path on the box's host cores (its own PyTorch code when
    oracle/_ref or /root/reference is there, else the oracle port), rank 0 only.  A STEP of this arm is one bounded sample."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import ref_import
    w = workload_cfg(args.workload or DEFAULT_WORKLOAD)
    budget = max(2.0, min(args.cpu_budget, 150.0 / max(1, args.steps + args.warmup)))
    sample = ref_sample if ref_import.available() else cpu_sample
    for _ in range(min(args.warmup, 1)):
        sample(w, args.iters, budget, args.zero_frac)
    vals, secs, wall = [], [], []
    for _ in range(args.steps):
        t0 = time.perf_counter()
        r = sample(w, args.iters, budget, args.zero_frac)
        wall.append(time.perf_counter() - t0)
        vals.append(r["value"]); secs.append(r["sample_seconds"])
    vals.sort()
    v = vals[len(vals) // 2]
    r["value"] = v
    cfg = config_dict(w, args)
    cfg["parallelism"] = f"CPU: {r['cores']} host threads, {'the reference PyTorch path (_knn_torch)' if r['kind'] == 'reference' else 'C/OpenMP port of the reference path'}"
    cfg["search"] = "brute force (torch.cdist + topk)" if r["kind"] == "reference" else "brute force"
    line = {"impl": "reference", "metric": "knn_corr_queries_per_s", "value": v, "unit": "queries/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": min(args.warmup, 1), "ms_per_step": 1e3 * sorted(wall)[len(wall) // 2],
            "step_is": "one bounded sample (1 frame, a track subsample); see cpu_baseline.sample",
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": cfg, "cpu_baseline": r,
            "e2e": {"value": v, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def config_dict(w, args, world=1, weak=False):
    return {"parallelism": "1 GPU: fused knn_corr kernel" if world == 1 else
            f"{world} GPUs: independent replicas, each rank runs the single-GPU path on its own sequence's window (no exchange)" if weak else
            (f"{world} GPUs: the FRAMES of the window (the batch axis of the reference's kNN correlation: independent units) are sharded round robin, frame t on rank "
             f"t % {world}; each rank builds and caches its frames' clouds of all views (nothing of a cloud crosses NVLink) and per iteration runs the single-GPU fused "
             f"kernel on its frames x all tracks, whose epilogue stores every result record into every rank's block (multimem.st over NVSwitch); one signal-pad barrier per "
             f"iteration, no NCCL on the data path" if getattr(args, "dist_mode", "") == "frames" else
             f"{world} GPUs: views sharded for the cloud build (features stay local); 16-byte point records all-gathered once per window; exact search per track shard; "
             f"index lists all-gathered; records of locally owned neighbours all-reduced (NCCL)" if getattr(args, "dist_mode", "tracks") == "tracks" else
             f"{world} GPUs: views sharded for the cloud build (features stay local, 16-byte point records multicast into every rank's replica); per iteration ONE persistent "
             f"kernel per rank: search of the rank's track shard -> 8-byte records into the OWNER's inbox (peer stores over NVLink) -> owner-side gather + correlation "
             f"-> result records multicast to every rank -> in-kernel epoch exchange; no NCCL, no barrier launch on the data path" if args.dist_mode == "fused" else
             f"{world} GPUs: views sharded for the cloud build (features stay local); point records, index lists and result records are stored by the producing kernels "
             f"straight into every rank's buffers (symmetric memory, NVSwitch multicast stores), two signal-pad barriers per iteration, no NCCL on the data path" if args.dist_mode == "p2p" else
             f"{world} GPUs: views sharded for the cloud build (features stay local), candidate keys all-to-all per track shard (two bounded phases), merged indices all-gathered, result records all-reduced (NCCL)"),"workload": f"{w['name']}: {w['V']} views {w['H']}x{w['W']} (stride {STRIDE} -> {w['Hs']}x{w['Ws']}), "
                        f"{w['V'] * w['Hs'] * w['Ws']} fused points/frame at level 0, window S={S_WINDOW}, {w['N']} tracks",
            "pipeline": ("window j+1's new frames are built on a side stream while window j's iterations run (every build inside the timed region)"
                         if world > 1 and not weak and getattr(args, "dist_mode", "") == "frames" and not getattr(args, "no_prefetch", False) else "none"),
            "levels": LEVELS, "k": K_NN, "iters": args.iters, "warm_start": bool(getattr(args, "warm_start", False)), "C": C_FEAT, "groups": 1, "search": getattr(args, "search", None),
            "queries_per_step": args.iters * LEVELS * S_WINDOW * w["N"], "zero_depth_frac": args.zero_frac,
            "step": ("one window advance of a sequence: 6 new frames built into the per-frame cloud cache (24 distinct frames, cycled: the new "
                     "frames differ every step) + iters x knn_corr over the 12-frame window" if getattr(args, "step", "advance") == "advance"
                     and not weak and getattr(args, "search", "raster") == "raster" else "one window, all 12 frames rebuilt every step"),
            "l2": "inputs_exceed_l2 (feature maps and clouds are > 126 MB and are rewritten every step)"}


# ------------------------------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------------------------------
def _peak():
    try:
        mp = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        return float(mp["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def run_ours(args):
    import torch
    import torch.distributed as dist
    from mvtracker_b200 import _lib, ops, synth
    from mvtracker_b200.dist import ShardedWindow, view_range
    from mvtracker_b200.hotpath import WindowClouds

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert world == args.gpus or world == 1, f"--gpus {args.gpus} but WORLD_SIZE={world}: launch with torchrun, one rank per GPU"
    assert args.gpus == 1 or world > 1, "for --gpus N > 1 launch with torchrun (one rank per GPU)"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    w = workload_cfg(args.workload or DEFAULT_WORKLOAD)
    fdtype = torch.float32 if args.dtype == "f32" else torch.bfloat16
    esz = 4 if args.dtype == "f32" else 2
    iters, N, V = args.iters, w["N"], w["V"]
    warmup = max(args.warmup, 3)
    L = _lib.lib()
    weak = args.scaling == "weak" and world > 1
    pw = 1 if weak else world          # ranks ONE window is sharded over

    # ---- synthetic inputs, identical on every rank (same seed); tracks are sampled from the full level-0 cloud ---------
    from mvtracker_b200.dist import FrameShardedSequence, ShardedSequence
    from mvtracker_b200.hotpath import WindowSequence
    advance = args.step == "advance" and not weak and args.search == "raster"
    frames_mode = pw > 1 and args.dist_mode == "frames"
    if frames_mode and (not advance or (4 * (S_WINDOW // 2)) % world):
        args.dist_mode, frames_mode = "p2p", False      # the frame pool (24 frames, cycled) must split evenly over the ranks
    HOP = S_WINDOW // 2
    S_POOL = 4 * HOP if advance else S_WINDOW      # advance: a pool of 24 distinct frames, cycled (window j = blocks j, j+1 mod 4)
    sc = synth.make_scene(V, w["H"], w["W"], S=S_POOL, C=C_FEAT, stride=STRIDE, zero_frac=args.zero_frac,
                          seed=72 + (rank if weak else 0), device=dev, dtype=fdtype)
    fm_pool, dp_pool = sc.fmaps[0], sc.depths[0, :, :, 0].contiguous()
    K_pool, E_pool = sc.intrs[0].contiguous(), sc.extrs[0].contiguous()
    c0 = ops.cloud_build(fm_pool, dp_pool, K_pool, E_pool, stride=STRIDE, levels=[0], want_fvec=False, want_spacing=False)[0]
    # tracks follow the surface at their pixel (+ N(0, 2 cm)), i.e. near the cloud in every frame like converged tracks
    pool_coords, _ = synth.make_tracks(c0.xyz, N, S_POOL, C=C_FEAT, seed=73, dtype=fdtype, valid=c0.xyz4[..., 3] > 0)
    _, base_targets = synth.make_tracks(c0.xyz[:S_WINDOW], N, S_WINDOW, C=C_FEAT, seed=73, dtype=fdtype)
    del c0

    def frames_of(j):          # pool frames of window j
        if not advance:
            return list(range(S_WINDOW))
        a, b = j % 4, (j + 1) % 4
        return list(range(HOP * a, HOP * a + HOP)) + list(range(HOP * b, HOP * b + HOP))

    g = torch.Generator(device=dev).manual_seed(74)
    jit = [0.01 * torch.randn(S_WINDOW, N, 3, device=dev, generator=g) for _ in range(iters)]
    n_types = 4 if advance else 1
    coords_win = [[(pool_coords[frames_of(t)] + jit[i]).contiguous() for i in range(iters)] for t in range(n_types)]
    cams_win = [(K_pool[:, frames_of(t)].contiguous(), E_pool[:, frames_of(t)].contiguous()) for t in range(n_types)]
    coords_it = coords_win[0]
    targets_it = [(base_targets * (1.0 - 0.1 * i)).contiguous() for i in range(iters)]
    width = LEVELS * K_NN * 4
    outs = [torch.empty(S_WINDOW, N, width, device=dev, dtype=torch.float32) for _ in range(iters)]
    v0, v1 = 0, V
    if frames_mode:        # this rank's frames of the pool (all views): pool frame p belongs to rank p % world
        fm_pool, dp_pool = fm_pool[:, rank::world].contiguous(), dp_pool[:, rank::world].contiguous()
        del sc
        torch.cuda.empty_cache()
    elif pw > 1:
        v0, v1 = view_range(rank, world, V)
        fm_pool, dp_pool = fm_pool[v0:v1].contiguous(), dp_pool[v0:v1].contiguous()
        del sc
        torch.cuda.empty_cache()
    cams_loc = [(Kw[v0:v1].contiguous(), Ew[v0:v1].contiguous()) for Kw, Ew in cams_win]
    fm, dp = fm_pool[:, :S_WINDOW], dp_pool[:, :S_WINDOW]          # window 0 (the whole input of --step window)
    K, E = cams_loc[0]
    K_all, E_all = cams_win[0]
    n_built = HOP if advance else S_WINDOW                         # frames whose clouds a steady-state step builds
    if frames_mode:
        n_built = (HOP + world - 1) // world                       # ... on the rank that builds most (all views of them)

    def local_run(ts):         # global frame numbers (consecutive multiples of world apart) -> slice of this rank's pool frames
        li = [((t % S_POOL) - rank) // world for t in ts]
        assert li == list(range(li[0], li[0] + len(li)))
        return slice(li[0], li[0] + len(li))

    # ---- one step = the hot path of one sliding window; output buffers are allocated once and rebuilt in place --------
    state = {"win": None, "seq": None, "j": 0}

    def step(ev=None, pipeline=True):
        if ev is not None: ev[0].record()
        if advance and frames_mode:
            # software pipeline: window j + 1's frames are built on a side stream while window j's iterations run (every build is
            # inside the timed region; --no-prefetch builds at the start of the step instead)
            def inputs(seq, jj):
                ts = seq.new_frames()
                new = local_run(ts) if ts else None
                return (fm_pool[:, new] if ts else None, dp_pool[:, new] if ts else None, cams_win[jj % 4][0], cams_win[jj % 4][1])
            if state["seq"] is None:
                state["seq"] = FrameShardedSequence(V, w["Hs"], w["Ws"], C_FEAT, fdtype, dev, S=S_WINDOW, k=K_NN)
                j = 0
                win = state["seq"].push(*inputs(state["seq"], 0))
            else:
                j = state["j"] + 1
                win = state["seq"].push() if state["seq"].has_prefetch else state["seq"].push(*inputs(state["seq"], j))
            if pipeline and not args.no_prefetch:
                state["seq"].prefetch(*inputs(state["seq"], j + 1))
            t = j % 4
            state["j"] = j
            cw = coords_win[t]
        elif advance:
            if state["seq"] is None:
                Hs, Ws = w["Hs"], w["Ws"]
                state["seq"] = (WindowSequence(V, Hs, Ws, C_FEAT, fdtype, dev, S=S_WINDOW, k=K_NN) if pw == 1 else
                                ShardedSequence(V, v0, v1, Hs, Ws, C_FEAT, fdtype, dev, S=S_WINDOW, k=K_NN, mode=args.dist_mode))
                j, new = 0, slice(0, S_WINDOW)
            else:
                j = state["j"] + 1
                b = (j + 1) % 4
                new = slice(HOP * b, HOP * b + HOP)
            t = j % 4
            if pw == 1:
                win = state["seq"].push(fm_pool[:, new], dp_pool[:, new], cams_win[t][0], cams_win[t][1])
                win.warm_start = bool(args.warm_start)
            else:
                win = state["seq"].push(fm_pool[:, new], dp_pool[:, new], cams_loc[t][0], cams_loc[t][1], cams_win[t][0], cams_win[t][1])
            state["j"] = j
            cw = coords_win[t]
        elif pw == 1:
            clouds = ops.cloud_build(fm, dp, K, E, stride=STRIDE, levels=range(LEVELS), want_spacing=(args.search == "grid"),
                                     reuse=None if state["win"] is None else state["win"].clouds,
                                     want_tiles=(args.search == "raster"))
            if ev is not None: ev[1].record()
            grids = None
            if args.search == "grid":
                grids = ops.build_grids([(c.xyz, None, c.spacing) for c in clouds], k_hint=K_NN, cell_scale=args.cell_scale)
            win = WindowClouds(clouds, grids, K_NN, 1, True, False)
            win.warm_start = bool(args.warm_start)
            cw = coords_win[0]
        else:
            win = ShardedWindow.build(fm.contiguous(), dp.contiguous(), K, E, v0, V, stride=STRIDE, levels=range(LEVELS), k=K_NN,
                                      reuse=state["win"], intrs_all=K_all, extrs_all=E_all, mode=args.dist_mode)
            cw = coords_win[0]
        if ev is not None and (advance or pw > 1): ev[1].record()
        if ev is not None: ev[2].record()
        for i in range(iters):
            # "p2p" / "fused": the result block is returned where the ranks' kernels stored it (symmetric memory; valid until the
            # next-but-one corr call), as the fused kernel writes into the caller's buffer at N = 1 -- no extra copy
            state["out"] = win.corr(targets_it[i], cw[i], out=None if getattr(win, "mode", "") in ("p2p", "fused", "frames") else outs[i])
            if ev is not None: ev[3 + i].record()
        state["win"], state["coords"] = win, cw

    def sync():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    for _ in range(warmup):
        step()
    sync()

    # ---- timed region: exactly K steps between two events on the launching (current) stream, barrier + sync around ---
    try:
        gpu_id = "GPU-" + str(torch.cuda.get_device_properties(local).uuid)
    except Exception:
        gpu_id = str(local)
    sampler = ClockSampler(gpu_id) if rank == 0 else None
    time.sleep(0.3)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n0 = L.mvt_launch_count()
    sync()
    t_wall0 = time.time()
    e0.record()
    for _ in range(args.steps):
        step()
    e1.record()
    sync()
    t_wall1 = time.time()
    launches = L.mvt_launch_count() - n0
    total_ms = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([total_ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms = float(t.item())
    ms_step = total_ms / args.steps
    queries = iters * LEVELS * S_WINDOW * N * (world if weak else 1)
    value = queries / (ms_step * 1e-3)

    # ---- per-kernel split (separate, untimed-for-the-metric pass with events between the stages) -------------------------
    n_split = max(3, min(args.steps, 10))
    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(3 + iters)] for _ in range(n_split)]
    sync()
    if frames_mode:
        step(pipeline=False)        # (adopts the window prefetched by the last timed step; the split below builds in line)
    built0 = state["seq"].frames_built if frames_mode else 0
    for e in evs:
        step(e, pipeline=False)
    if frames_mode:                 # frames THIS rank built per split step (a rank builds 0 or 1 frame per advance at N = 8)
        n_built = max((state["seq"].frames_built - built0) / n_split, 1e-9)
    # keep the GPU busy long enough for nvidia-smi to have samples of the loaded clocks
    # (the SAME number of extra steps on every rank -- from the all-reduced time, not from each rank's own wall clock: the
    # sharded modes synchronise inside step(), ranks that enqueue different numbers of steps would wait for each other forever)
    for _ in range(max(0, min(2000, int((1200.0 - total_ms) / max(ms_step, 1e-3)) + 1))):
        step()
    sync()
    t_cloud = sum(e[0].elapsed_time(e[1]) for e in evs) / n_split
    overlapped = frames_mode and not args.no_prefetch
    t_grid = sum(e[1].elapsed_time(e[2]) for e in evs) / n_split
    t_corr = sum(e[2 + i].elapsed_time(e[3 + i]) for e in evs for i in range(iters)) / (n_split * iters)
    clocks = sampler.window(t_wall0, time.time()) if sampler else None

    # ---- roofline (algorithmic bytes / measured time / measured HBM peak), per GPU -----------------------------------------
    peak, peak_src = _peak()
    v_loc = fm.shape[0]        # (frames mode: all views, of ceil(HOP / world) frames)
    q_bytes = 12 + C_FEAT * esz + K_NN * (C_FEAT * esz + 12) + K_NN * 4 * 4          # 9164 B (fp32) / 4812 B (bf16)
    corr_bytes = LEVELS * S_WINDOW * N * q_bytes / pw                                 # per rank: its share of the gathers
    pts = sum(v_loc * (w["Hs"] >> l) * (w["Ws"] >> l) for l in range(LEVELS)) * n_built
    cloud_bytes = int((fm[:, :1].numel() * esz + dp[:, :1].numel() * 4) * n_built + pts * (C_FEAT * esz + 12))   # SURVEY.md 8d (xyz4 records not counted)
    kern = {
        "knn_corr": {"ms": t_corr, "launches_per_step": iters, "alg_bytes": corr_bytes, "gbs": corr_bytes / t_corr / 1e6,
                     "what": "fused knn_corr kernel" if pw == 1 else
                     "fused knn_corr kernel on this rank's frames, records stored into every rank's block + one barrier (per iteration)" if frames_mode else
                     "search kernel + index exchange + corr_from_idx + result exchange (per iteration)"},
        "cloud_build": {"ms": t_cloud, "launches_per_step": 1, "alg_bytes": cloud_bytes, "gbs": cloud_bytes / t_cloud / 1e6, "frames_built": n_built,
                        "what": "camera_prep + cloud_build (4 levels, one pass) + tile depth tables" + ("" if pw == 1 or args.dist_mode != "tracks" else " + all-gather of the 16-byte records")
                                + (" -- timed here in line; in the timed steps it runs on a side stream under the previous window's iterations, so ms_per_step < the sum of the parts" if overlapped else "")},
    }
    if args.search == "grid" and pw == 1:
        kern["grid_build"] = {"ms": t_grid, "launches_per_step": 6, "alg_bytes": 0, "gbs": 0.0}
    step_sum = sum(kv["ms"] * kv["launches_per_step"] if name != "grid_build" else kv["ms"] for name, kv in kern.items())
    for name, kv in kern.items():
        kv["frac"] = kv["gbs"] / peak
        kv["share_of_step"] = (kv["ms"] * kv["launches_per_step"] if name != "grid_build" else kv["ms"]) / step_sum
    dom = "knn_corr" if t_corr * iters >= t_cloud else "cloud_build"
    roofline = {"bound": "hbm", "kernel": dom, "achieved": kern[dom]["gbs"], "peak": peak, "peak_source": peak_src,
                "unit": "GB/s", "frac": kern[dom]["gbs"] / peak, "traffic": ncu_traffic(w["name"], args.dtype, dom, world),
                "alg_bytes_per_launch": kern[dom]["alg_bytes"], "ms_per_launch": kern[dom]["ms"],
                "step_frac": (corr_bytes * iters + cloud_bytes) / (ms_step * 1e-3) / 1e9 / peak}

    line = {"metric": "knn_corr_queries_per_s", "value": value, "unit": "queries/s", "n_gpus": world, "steps": args.steps,
            "warmup": warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak" if weak else "strong",
            "vs_baseline": None, "dtype": args.dtype, "data": "synthetic", "config": config_dict(w, args, world, weak), "clocks": clocks,
            "gpu_launches": int(launches), "roofline": roofline, "kernels": kern}

    # BASELINE.json's second metric, for the part of MVTracker.forward this repo replaces: the sequence's T frames over the
    # hot-path time of its sliding windows (encoder / transformer are stock PyTorch in the reference and not timed here)
    n_win = len(range(0, max(w["T"] - S_WINDOW // 2, 1), S_WINDOW // 2))
    line["hot_path_frames_per_s"] = {"value": w["T"] / (n_win * ms_step * 1e-3), "frames": w["T"], "windows": n_win,
                                     "what": "T / (windows x step): hot path only, not MVTracker.forward end to end"}

    # ---- the build of a whole 12-frame window (what the reference does every window), for the build kernel's own roofline --------
    if advance and pw == 1:
        full_bytes = cloud_bytes * S_WINDOW // n_built
        ws = None
        for _ in range(2):
            ws = ops.cloud_build(fm, dp, K, E, stride=STRIDE, levels=range(LEVELS), want_spacing=False, want_tiles=True, reuse=ws)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(3):
            ops.cloud_build(fm, dp, K, E, stride=STRIDE, levels=range(LEVELS), want_spacing=False, want_tiles=True, reuse=ws)
        b.record(); torch.cuda.synchronize()
        t_full = a.elapsed_time(b) / 3
        kern["cloud_build_full"] = {"ms": t_full, "alg_bytes": full_bytes, "gbs": full_bytes / t_full / 1e6, "frac": full_bytes / t_full / 1e6 / peak,
                                    "frames_built": S_WINDOW, "what": "all 12 frames of a window rebuilt (not part of the step)"}
        del ws

    # ---- the timed step's output, sampled against the oracle (every N) ----------------------------------------------------------
    if not args.no_verify and not weak:
        line.update(verify_output(state["win"], state["out"], targets_it[iters - 1], state["coords"][iters - 1], world, rank,
                                  args.dtype == "bf16"))

    # ---- two more track distributions on the same window (cold search each), N = 1 ---------------------------------------------
    if world == 1 and not args.no_distributions and args.search == "raster":
        fr = frames_of(state["j"])
        kern.update(time_distributions(args, w, ops, synth, WindowClouds, state["win"], fm_pool[:, fr].contiguous(), dp_pool[:, fr].contiguous(),
                                       cams_win[state["j"] % n_types][0], cams_win[state["j"] % n_types][1], targets_it[0], dev, peak, esz))

    # ---- adjacent kernels, timed outside the step: not part of `value` -----------------------------
    if world == 1 and not args.no_adjacent and args.search == "raster":
        line["adjacent_kernels"] = time_adjacent(ops, state["win"], targets_it[0], coords_it[0], w, esz, peak, dev)

    # ---- e2e: same metric through the reference-facing API with HOST buffers ---------------------------------------------
    if not args.no_e2e:
        e2e = run_e2e(args, w, world, rank, dev, fm_pool, dp_pool, cams_loc, cams_win, coords_win, targets_it, v0, v1, advance, weak)
        line["e2e"] = e2e
    state["win"] = None
    state["out"] = None
    if sampler:
        sampler.stop()
    if world == 1 and not args.no_model:
        # BASELINE.json metric (ii): MVTracker.forward end to end, reference seam vs drop-in, on this GPU
        del fm, dp, fm_pool, dp_pool, sc, outs, coords_it, coords_win, targets_it, pool_coords, base_targets
        state["seq"] = None
        torch.cuda.empty_cache()
        try:
            line["e2e_frames_per_s"] = model_frames_per_s([n for n in args.model_workloads.split(",") if n], dev, iters)
        except Exception as e:      # the legs use the reference's model code; a failure there must not void the bench line
            line["e2e_frames_per_s"] = {"error": f"{type(e).__name__}: {e}"[:300]}
    if rank == 0:
        if world == 1 and not args.no_cpu:
            from oracle import ref_import
            line["cpu_baseline"] = (ref_sample if ref_import.available() else cpu_sample)(w, iters, args.cpu_budget, args.zero_frac)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def time_adjacent(ops, win, targets, coords, w, esz, peak, dev, reps=5):
    """track_input (row f1) on the window's [N, S, 581] buffer; knn_corr_backward + cloud_build_backward (row f2) at level 0.
    CUDA events on the current stream, 2 warm-up + `reps` timed launches each; algorithmic bytes as in DESIGN.md."""
    import torch
    N, S, C, E, V, Hs, Ws = w["N"], S_WINDOW, C_FEAT, 64, w["V"], w["Hs"], w["Ws"]
    width = LEVELS * K_NN * 4
    D = 3 * E + 3 + width + C + 2

    def timed(fn, prep=None):
        for _ in range(2):
            if prep: prep()
            fn()
        ms = 0.0
        for _ in range(reps):
            if prep: prep()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); fn(); b.record()
            torch.cuda.synchronize()
            ms += a.elapsed_time(b)
        return ms / reps

    out = {}
    g = torch.Generator(device=dev).manual_seed(75)
    x = torch.empty(N, S, D, device=dev)
    pos, tim = torch.randn(N, D, device=dev, generator=g), torch.randn(S, D, device=dev, generator=g)
    mask, vis = torch.ones(S, N, device=dev), torch.rand(S, N, device=dev, generator=g)
    win.corr(targets, coords, out=ops.corr_columns(x, E, width))
    ms = timed(lambda: ops.track_input(coords, targets, mask, vis, pos, tim, E, width, out=x, corr_in_out=True))
    b = N * S * (D * 4 + width * 4 + C * esz + 12 + 8) + (N + S) * D * 4
    out["track_input"] = {"ms": ms, "alg_bytes": b, "gbs": b / ms / 1e6, "frac": b / ms / 1e6 / peak,
                          "what": f"[{N},{S},{D}] transformer input rows written once (corr columns + pos + time in place)"}
    c0 = win.clouds[0]
    _, idx = ops.knn_corr([(c0, c0.fvec, K_NN)], targets, coords, return_idx=True)
    gout = torch.randn(S, N, K_NN, 4, device=dev, generator=g)
    gfv = torch.zeros(c0.fvec.shape, device=dev, dtype=torch.float32)
    gtg = torch.zeros(S, N, C, device=dev)
    ms = timed(lambda: ops.knn_corr_backward(c0.fvec, targets, idx[0], gout, grad_targets=gtg, grad_fvec=gfv))
    b = S * N * (C * esz + K_NN * C * esz + K_NN * C * 4 + C * 4 + K_NN * 4 + K_NN * 4)
    out["knn_corr_backward"] = {"ms": ms, "alg_bytes": b, "gbs": b / ms / 1e6, "frac": b / ms / 1e6 / peak,
                                "what": "level 0: gather of the neighbour rows + one 16-byte vector reduction per lane and neighbour"}
    gfm = torch.empty(V, S, C, Hs, Ws, device=dev)
    ms = timed(lambda: ops.cloud_build_backward(gfv, V, Hs, Ws, 0, grad_fmaps=gfm, accumulate=False))
    b = 2 * gfm.numel() * 4                        # read grad_fvec, write grad_fmaps
    out["cloud_build_backward"] = {"ms": ms, "alg_bytes": b, "gbs": b / ms / 1e6, "frac": b / ms / 1e6 / peak,
                                   "what": "level 0 un-pool + transpose (point-major rows -> channel-major planes)"}
    return out


def verify_output(win, out, targets, coords, world, rank, bf16, n_tracks=64, frames=(0, S_WINDOW - 1)):
    """Samples the LAST timed step's output: n_tracks tracks x `frames` x every level against oracle.corr_sample (brute
    force, canonical (d2, index) order) on the clouds the step built.  Offsets must be bit-equal (same neighbour sets in the
    same order); correlations within 1e-4 (fp32) / 2e-2 (bf16) of the level's largest |corr|.  N > 1: every rank checks the
    replicated result block; the feature rows of the sampled frames are all-gathered for the check only."""
    import numpy as np
    import torch
    import torch.distributed as dist
    from oracle import oracle
    S, N, _ = out.shape
    sel = torch.linspace(0, N - 1, min(n_tracks, N), device=out.device).long()
    worst, bad_sets, checked = 0.0, 0, 0
    sharded = hasattr(win, "full") and getattr(win, "full", None) is not None
    by_frames = getattr(win, "mode", "") == "frames"       # every rank checks the frames IT owns, in the replicated block
    n_levels = len(win.levels) if by_frames else len(win.full) if sharded else len(win.clouds)
    if by_frames:
        frames = sorted({0, len(win.run) - 1}) if win.run else ()
    for l in range(n_levels):
        for f in frames:
            if by_frames:
                xyz, fvec = win.clouds[l].xyz[f], win.clouds[l].fvec[f]
                f = win.pos0 + f * win.world                # local frame -> position in the window
            elif sharded:
                xyz = win.full[l].xyz4[f, :, :3].contiguous()
                mine = win.clouds[l].fvec[f].contiguous()
                parts = [torch.empty_like(mine) for _ in range(world)]
                dist.all_gather(parts, mine)
                fvec = torch.cat(parts, 0)
            else:
                xyz, fvec = win.clouds[l].xyz[f], win.clouds[l].fvec[f]
            ref = oracle.corr_sample(K_NN, xyz[None].cpu().numpy(), fvec[None].float().cpu().numpy(),
                                     targets[f:f + 1, sel].float().cpu().numpy(), coords[f:f + 1, sel].cpu().numpy())[0]
            got = out[f, sel, l * K_NN * 4:(l + 1) * K_NN * 4].reshape(-1, K_NN, 4).cpu().numpy()
            bad_sets += int((got[..., 1:] != ref[..., 1:]).any(-1).any(-1).sum())
            err = float(np.abs(got[..., 0] - ref[..., 0]).max() / max(1.0, float(np.abs(ref[..., 0]).max())))
            worst = max(worst, err)
            checked += got.shape[0]
    tol = 2e-2 if bf16 else 1e-4
    ok = bad_sets == 0 and worst <= tol
    if world > 1:
        t = torch.tensor([float(bad_sets), worst, float(not ok)], device=out.device, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        bad_sets, worst, ok = int(t[0].item()), float(t[1].item()), t[2].item() == 0.0
    return {"verified": bool(ok), "queries_checked": checked, "neighbour_set_mismatches": bad_sets, "max_corr_err": worst,
            "tolerance": tol, "against": "oracle.corr_sample (brute force) on the clouds of the last timed step, every rank" if world > 1
            else "oracle.corr_sample (brute force) on the clouds of the last timed step"}


def time_distributions(args, w, ops, synth, WindowClouds, win, fm, dp, K, E, base_targets, dev, peak, esz, reps=5):
    """The raster search's cost depends on where the tracks are (SURVEY.md 8d asks 0 / 5 / 40 % invalid depth; the reference
    starts every window with the SAME 3-D point in all S frames, mvtracker.py:510).  Times one fused knn_corr launch (cold
    search) for two more distributions on the benched window: `static` = frame 0's surface point repeated in all frames
    (+ N(0, 2 cm)): tracks of moving objects end up in free space in later frames; `zero40` = the same scene with 40 % of the
    depth pixels invalid.  Not part of `value`."""
    import torch
    N = w["N"]
    q_bytes = 12 + C_FEAT * esz + K_NN * (C_FEAT * esz + 12) + K_NN * 4 * 4
    alg = LEVELS * S_WINDOW * N * q_bytes
    out_buf = torch.empty(S_WINDOW, N, LEVELS * K_NN * 4, device=dev, dtype=torch.float32)

    def timed(wn, coords):
        for _ in range(10):         # steady state of THIS distribution: ops.knn_corr's deferral policy samples every 4th launch
            wn.corr(base_targets, coords, out=out_buf)      # and acts one call later (no host sync in the product path)
            torch.cuda.synchronize()
        ms = 0.0
        for _ in range(reps):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(); wn.corr(base_targets, coords, out=out_buf); b.record()
            torch.cuda.synchronize()
            ms += a.elapsed_time(b)
        ms /= reps
        pol = ops.hard_query_policy(dev)
        return {"ms": ms, "gbs": alg / ms / 1e6, "frac": alg / ms / 1e6 / peak, "deferral": bool(pol.deferring),
                "hard_query_fraction": pol.last_fraction}

    res = {}
    c0 = win.clouds[0]
    valid0 = c0.xyz4[0, :, 3] > 0
    static_coords, _ = synth.make_tracks(c0.xyz[0], N, S_WINDOW, C=C_FEAT, seed=76, valid=valid0)
    res["knn_corr_static"] = dict(timed(win, static_coords), what="same 3-D point in all 12 frames (frame 0's surface + N(0, 2 cm)), "
                                  "the reference's window initialisation (mvtracker.py:510)")
    # 40 % invalid depth: rebuild the window's clouds in place from the same feature maps
    g = torch.Generator(device=dev).manual_seed(77)
    dp40 = torch.where(torch.rand(dp.shape, device=dev, generator=g) < (0.4 - args.zero_frac) / max(1e-6, 1.0 - args.zero_frac),
                       torch.zeros_like(dp), dp)
    clouds = ops.cloud_build(fm, dp40, K, E, stride=STRIDE, levels=range(LEVELS), want_spacing=False, reuse=win.clouds, want_tiles=True)
    w40 = WindowClouds(clouds, None, K_NN, 1, True, False)
    coords40, _ = synth.make_tracks(clouds[0].xyz, N, S_WINDOW, C=C_FEAT, seed=73, valid=clouds[0].xyz4[..., 3] > 0)
    res["knn_corr_zero40"] = dict(timed(w40, coords40), what="tracks follow the surface, 40 % of the depth pixels invalid (0)")
    ops.cloud_build(fm, dp, K, E, stride=STRIDE, levels=range(LEVELS), want_spacing=False, reuse=win.clouds, want_tiles=True)     # restore
    return res


def model_frames_per_s(names, dev, iters):
    """BASELINE.json metric (ii): end-to-end tracked frames/s = T / wall time of the reference's own MVTracker.forward
    (evaluator_3dpt.py:514-521), random-init weights of configs/model/mvtracker.yaml, host inputs copied to the device
    inside the timed region -- once with the reference's stock seam (its torch kNN / gather / einsum path ON THIS B200: the
    GPU bar BASELINE.md names) and once with mvtracker_b200.dropin.install().  Needs oracle/_ref (oracle/make_ref.py)."""
    import torch
    from mvtracker_b200 import synth
    from oracle import model_harness as mh
    from oracle import ref_import
    if not ref_import.available():
        return {"unavailable": "reference model not staged: run python oracle/make_ref.py where /root/reference exists"}
    out = {}
    model = mh.build_model(seed=72, device=dev, use_flash_attention=True)
    for name in names:
        wl = synth.WORKLOADS[name]
        seq_host = mh.make_sequence(wl["V"], wl["T"], wl["H"], wl["W"], wl["N"], seed=72, device="cpu", staggered=False)
        pinned = {k: getattr(seq_host, k).pin_memory() for k in ("rgbs", "depths", "intrs", "extrs", "query_points")}
        res = {"frames": wl["T"], "views": wl["V"], "tracks": wl["N"], "resolution": [wl["H"], wl["W"]], "iters": iters}
        # _knn_torch materialises [S, N, P0] fp32 per level-0 call; skip the stock arm where that cannot fit
        need = 4 * 12 * wl["N"] * wl["V"] * (wl["H"] // STRIDE) * (wl["W"] // STRIDE) * 2.5
        arms = ["dropin"] + (["stock"] if need < 0.6 * torch.cuda.mem_get_info(dev)[0] else [])
        outs = {}
        for seam in arms:
            best = None
            for rep in range(2):                                    # first repetition = warm-up (cuDNN autotune, allocator)
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                seq = type(seq_host)(**{**vars(seq_host), **{k: v.to(dev, non_blocking=True) for k, v in pinned.items()}})
                o, _ = mh.forward(model, seq, iters=iters, seam=seam)
                traj = o["traj_e"].cpu()
                dt = time.perf_counter() - t0
                best = dt if best is None or rep > 0 else best
            outs[seam] = traj
            res[seam] = {"frames_per_s": wl["T"] / best, "seconds": best}
            del seq, o
            torch.cuda.empty_cache()
        if "stock" in res:
            res["speedup"] = res["dropin"]["frames_per_s"] / res["stock"]["frames_per_s"]
            res["max_traj_diff"] = float((outs["dropin"] - outs["stock"]).abs().max())
        else:
            res["stock"] = {"unavailable": "_knn_torch's [S, N, P] distance tensor does not fit this GPU at this config"}
        out[name] = res
    return out


def run_e2e(args, w, world, rank, dev, fm_pool, dp_pool, cams_loc, cams_win, coords_win, targets_it, v0, v1, advance, weak=False):
    """The same step through the repo's public API with HOST buffers: every step copies the frames the window adds (6 with
    --step advance, 12 with --step window; this rank's views) and the window's cameras from pinned host memory, advances the
    sequence (N = 1: mvtracker_b200.hotpath.WindowSequence / WindowClouds; N > 1: mvtracker_b200.dist.ShardedSequence /
    ShardedWindow), and per iteration copies track positions + features in and the correlation block out (N > 1: each
    rank reads back the rows of ITS track shard -- together the ranks return the block once)."""
    import torch
    import torch.distributed as dist
    from mvtracker_b200.dist import FrameShardedSequence, ShardedSequence, ShardedWindow, track_shard
    from mvtracker_b200.hotpath import WindowClouds, WindowSequence
    iters, N, V = args.iters, w["N"], w["V"]
    pw = 1 if weak else world
    HOP = S_WINDOW // 2
    if pw > 1 and args.dist_mode == "frames" and advance:
        return run_e2e_frames(args, w, world, rank, dev, fm_pool, dp_pool, cams_win, coords_win, targets_it)
    nblk = fm_pool.shape[1] // HOP
    host_fm = [fm_pool[:, HOP * b:HOP * b + HOP].contiguous().cpu().pin_memory() for b in range(nblk)]
    host_dp = [dp_pool[:, HOP * b:HOP * b + HOP].contiguous().cpu().pin_memory() for b in range(nblk)]
    host_cam = [[t.cpu().pin_memory() for t in (cl[0], cl[1], cw[0], cw[1])] for cl, cw in zip(cams_loc, cams_win)]
    host_ct = [[(c.cpu().pin_memory(), t.cpu().pin_memory()) for c, t in zip(cw, targets_it)] for cw in coords_win]
    Ms = track_shard(pw, N)
    lo, hi = (rank * Ms, min((rank + 1) * Ms, N)) if pw > 1 else (0, N)
    host_out = [torch.empty(S_WINDOW, max(hi - lo, 0), LEVELS * K_NN * 4, dtype=torch.float32).pin_memory() for _ in range(iters)]
    n_new = HOP if advance else S_WINDOW
    h2d = (host_fm[0].numel() * host_fm[0].element_size() + host_dp[0].numel() * 4) * (n_new // HOP) \
        + sum(t.numel() * t.element_size() for t in host_cam[0]) \
        + sum(c.numel() * c.element_size() + t.numel() * t.element_size() for c, t in host_ct[0])
    d2h = sum(t.numel() * t.element_size() for t in host_out)
    state = {"seq": None, "win": None, "j": 0}
    Hs, Ws = w["Hs"], w["Ws"]
    fdtype = fm_pool.dtype

    def step():
        first = state["seq"] is None and state["win"] is None
        if advance:
            j = 0 if first else state["j"] + 1
            blocks = [0, 1] if first else [(j + 1) % 4]
        else:
            j, blocks = 0, [0, 1]
        t = j % len(host_cam)
        fmd = torch.cat([host_fm[b].to(dev, non_blocking=True) for b in blocks], 1) if len(blocks) > 1 else host_fm[blocks[0]].to(dev, non_blocking=True)
        dpd = torch.cat([host_dp[b].to(dev, non_blocking=True) for b in blocks], 1) if len(blocks) > 1 else host_dp[blocks[0]].to(dev, non_blocking=True)
        Kl, El, Ka, Ea = [x.to(dev, non_blocking=True) for x in host_cam[t]]
        if advance:
            if first:
                state["seq"] = (WindowSequence(V, Hs, Ws, C_FEAT, fdtype, dev, S=S_WINDOW, k=K_NN) if pw == 1 else
                                ShardedSequence(V, v0, v1, Hs, Ws, C_FEAT, fdtype, dev, S=S_WINDOW, k=K_NN, mode=args.dist_mode))
            win = state["seq"].push(fmd, dpd, Ka, Ea) if pw == 1 else state["seq"].push(fmd, dpd, Kl, El, Ka, Ea)
        elif pw == 1:
            win = WindowClouds.build(fmd, dpd, Ka, Ea, stride=STRIDE, k=K_NN)
        else:
            win = ShardedWindow.build(fmd, dpd, Kl, El, v0, V, stride=STRIDE, levels=range(LEVELS), k=K_NN, reuse=state["win"],
                                      intrs_all=Ka, extrs_all=Ea, mode=args.dist_mode)
        state["win"], state["j"] = win, j
        for i in range(iters):
            coords = host_ct[t][i][0].to(dev, non_blocking=True)
            ffeats = host_ct[t][i][1].to(dev, non_blocking=True)
            res = win.corr(ffeats, coords)
            if hi > lo:
                host_out[i].copy_(res[:, lo:hi], non_blocking=True)

    def sync():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    n = max(3, min(args.steps, 10))
    for _ in range(3):
        step()
    sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        step()
    e1.record()
    sync()
    ms = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
        hb = torch.tensor([float(h2d), float(d2h)], device=dev, dtype=torch.float64)
        dist.all_reduce(hb)
        h2d, d2h = int(hb[0].item()), int(hb[1].item())
    ms /= n
    state["seq"] = state["win"] = None
    queries = iters * LEVELS * S_WINDOW * N * (world if weak else 1)
    return {"value": queries / (ms * 1e-3), "unit": "queries/s", "ms_per_step": ms, "h2d_bytes_per_step": int(h2d),
            "d2h_bytes_per_step": int(d2h), "steps": n,
            "api": ("mvtracker_b200.hotpath." + ("WindowSequence.push" if advance else "WindowClouds.build") + " / .corr" if pw == 1 else
                    "mvtracker_b200.dist." + ("ShardedSequence.push" if advance else "ShardedWindow.build") + " / .corr")
                   + ", pinned host buffers, all ranks summed"}


def run_e2e_frames(args, w, world, rank, dev, fm_loc, dp_loc, cams_win, coords_win, targets_it):
    """run_e2e for --dist-mode frames: every step copies THIS rank's new frames (all views) and the window's cameras from pinned
    host memory and advances mvtracker_b200.dist.FrameShardedSequence; per iteration track positions + features go in and each
    rank reads back the rows of the frames it computed (together the ranks return the block once)."""
    import torch
    import torch.distributed as dist
    from mvtracker_b200.dist import FrameShardedSequence
    iters, N, V = args.iters, w["N"], w["V"]
    S_POOL = fm_loc.shape[1] * world
    host_fm, host_dp = fm_loc.cpu().pin_memory(), dp_loc.cpu().pin_memory()            # [V, S_POOL / world, ...]
    host_cam = [[t.cpu().pin_memory() for t in cw] for cw in cams_win]
    host_ct = [[(c.cpu().pin_memory(), t.cpu().pin_memory()) for c, t in zip(cw, targets_it)] for cw in coords_win]
    most = (S_WINDOW + world - 1) // world
    host_out = [torch.empty(most, N, LEVELS * K_NN * 4, dtype=torch.float32).pin_memory() for _ in range(iters)]
    state = {"seq": None, "j": 0, "h2d": 0, "d2h": 0}

    def step(count=False):
        if state["seq"] is None:
            state["seq"], j = FrameShardedSequence(V, w["Hs"], w["Ws"], C_FEAT, fm_loc.dtype, dev, S=S_WINDOW, k=K_NN), 0
        else:
            j = state["j"] + 1
        seq, t = state["seq"], j % len(host_cam)
        ts = seq.new_frames()
        fmd = dpd = None
        if ts:
            a = ((ts[0] % S_POOL) - rank) // world
            fmd = torch.empty((V, len(ts)) + tuple(host_fm.shape[2:]), device=dev, dtype=host_fm.dtype)
            dpd = torch.empty((V, len(ts)) + tuple(host_dp.shape[2:]), device=dev, dtype=host_dp.dtype)
            for v in range(V):      # one contiguous pinned run per view (a strided slice of pinned memory would be staged by the host)
                fmd[v].copy_(host_fm[v, a:a + len(ts)], non_blocking=True)
                dpd[v].copy_(host_dp[v, a:a + len(ts)], non_blocking=True)
        Ka, Ea = [x.to(dev, non_blocking=True) for x in host_cam[t]]
        win = seq.push(fmd, dpd, Ka, Ea)
        state["j"] = j
        nb = (0 if fmd is None else fmd.numel() * fmd.element_size() + dpd.numel() * 4) + Ka.numel() * 4 + Ea.numel() * 4
        nd = 0
        for i in range(iters):
            coords = host_ct[t][i][0].to(dev, non_blocking=True)
            ffeats = host_ct[t][i][1].to(dev, non_blocking=True)
            res = win.corr(ffeats, coords)
            nw = len(win.run)
            if nw:
                host_out[i][:nw].copy_(res[win.pos0::world][:nw], non_blocking=True)
            nb += coords.numel() * 4 + ffeats.numel() * ffeats.element_size()
            nd += nw * N * LEVELS * K_NN * 16
        if count:
            state["h2d"] += nb
            state["d2h"] += nd

    def sync():
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()

    n = max(4, min(args.steps, 12))
    for _ in range(3):
        step()
    sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        step(True)
    e1.record()
    sync()
    tt = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
    dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    hb = torch.tensor([state["h2d"] / n, state["d2h"] / n], device=dev, dtype=torch.float64)
    dist.all_reduce(hb)
    ms = float(tt.item()) / n
    state["seq"] = None
    queries = iters * LEVELS * S_WINDOW * N
    return {"value": queries / (ms * 1e-3), "unit": "queries/s", "ms_per_step": ms, "h2d_bytes_per_step": int(hb[0].item()),
            "d2h_bytes_per_step": int(hb[1].item()), "steps": n,
            "api": "mvtracker_b200.dist.FrameShardedSequence.push / .corr, pinned host buffers, all ranks summed (mean per step)"}


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
