"""A/B timing of libmvt builds: python tools/ab.py build/libmvt_a.so build/libmvt_b.so ... [-- bench args]
Runs bench.py (kernel timings only) once per library and workload and prints knn_corr / cloud_build ms per launch."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
args = sys.argv[1:]
extra = []
if "--" in args:
    i = args.index("--")
    args, extra = args[:i], args[i + 1:]
workloads = os.environ.get("AB_WORKLOADS", "panoptic,droid").split(",")
for lib in args:
    for wl in workloads:
        env = dict(os.environ, MVT_LIB=os.path.join(ROOT, lib))
        out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--no-cpu", "--no-e2e", "--steps", "10", "--warmup", "3",
                              "--workload", wl] + extra, env=env, capture_output=True, text=True)
        try:
            line = json.loads(out.stdout.strip().splitlines()[-1])
            ks = line["kernels"]
            print(f"{lib:28s} {wl:9s} knn_corr {ks['knn_corr']['ms']:.4f} ms  frac {ks['knn_corr']['frac']:.3f}   "
                  f"cloud_build {ks['cloud_build']['ms']:.3f} ms  step {line['ms_per_step']:.3f} ms", flush=True)
        except Exception as e:  # noqa: BLE001
            print(lib, wl, "FAILED", e, out.stderr[-400:], flush=True)
