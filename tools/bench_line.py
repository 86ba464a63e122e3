"""Pretty-print the last JSON line of a bench.py run read from stdin: python bench.py ... | python tools/bench_line.py [tag]"""
import json, sys
lines = [l for l in sys.stdin.read().splitlines() if l.startswith("{")]
d = json.loads(lines[-1])
k = d.get("kernels", {})
print(" ".join(sys.argv[1:]), d["config"]["workload"].split(":")[0], d.get("dtype"), f"N={d['n_gpus']}", d.get("scaling"),
      f"value {d['value']:.4g}", f"ms/step {d['ms_per_step']:.3f}", {n: (round(v["ms"], 3), round(v["frac"], 3)) for n, v in k.items()},
      "launches", d.get("gpu_launches"), "e2e", d.get("e2e", {}).get("value"))
