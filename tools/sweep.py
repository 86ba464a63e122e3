"""BASELINE.json config 5 (scaling sweep) on the GPUs of this launch: views x level-0 cloud points x tracks.
    python tools/sweep.py [--gpus N] [--full]      -> one line per point, and a markdown table at the end
Each point is one `bench.py --workload sweep:V=..,P0=..,N=..` run (device-timed, no CPU / e2e legs)."""
import argparse, json, os, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ap = argparse.ArgumentParser()
ap.add_argument("--gpus", type=int, default=1)
ap.add_argument("--full", action="store_true", help="the whole 4 x 4 x 4 grid (default: a 13-point cross through it)")
a = ap.parse_args()
Vs, Ps, Ns = [2, 4, 8, 16], ["128k", "512k", "1M", "4M"], ["1k", "4k", "16k", "64k"]
if a.full:
    pts = [(v, p, n) for v in Vs for p in Ps for n in Ns]
else:   # vary one axis at a time around (8 views, 1M points, 4k tracks) + the two corners
    pts = sorted(set([(v, "1M", "4k") for v in Vs] + [(8, p, "4k") for p in Ps] + [(8, "1M", n) for n in Ns]
                     + [(2, "128k", "1k"), (16, "4M", "64k")]), key=lambda t: (t[0], Ps.index(t[1]), Ns.index(t[2])))
rows = []
for i, (v, p, n) in enumerate(pts):
    wl = f"sweep:V={v},P0={p},N={n}"
    cmd = [sys.executable]
    if a.gpus > 1:
        cmd += ["-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={a.gpus}", "--master-addr", "127.0.0.1",
                "--master-port", str(29700 + i)]
    cmd += [os.path.join(ROOT, "bench.py"), "--gpus", str(a.gpus), "--workload", wl, "--steps", "5", "--warmup", "3", "--no-cpu", "--no-e2e"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    line = [l for l in r.stdout.splitlines() if l.startswith("{")]
    if r.returncode != 0 or not line:
        print(wl, "FAILED", r.stderr[-300:], flush=True)
        continue
    d = json.loads(line[-1])
    k = d["kernels"]
    rows.append((v, p, n, d["ms_per_step"], k["cloud_build"]["ms"], k["cloud_build"]["frac"], k["knn_corr"]["ms"], k["knn_corr"]["frac"], d["value"]))
    print(wl, f"step {d['ms_per_step']:.3f} ms  cloud_build {k['cloud_build']['ms']:.3f} ms ({k['cloud_build']['frac']:.2f})  "
              f"knn_corr {k['knn_corr']['ms']:.3f} ms ({k['knn_corr']['frac']:.2f})  {d['value']:.3g} queries/s", flush=True)
print(f"\n| views | level-0 points / frame | tracks | step (ms) | cloud_build (ms, frac) | knn_corr per iteration (ms, frac) | queries/s ({a.gpus} GPU) |")
print("|---|---|---|---|---|---|---|")
for v, p, n, ms, tc, fc, tk, fk, val in rows:
    print(f"| {v} | {p} | {n} | {ms:.2f} | {tc:.3f}, {fc:.2f} | {tk:.3f}, {fk:.2f} | {val:.3g} |")
