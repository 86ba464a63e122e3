"""Pretty-print the last JSON line of a bench.py run: python tools/bench_line.py FILE [tag]   (or: ... | python tools/bench_line.py - [tag])"""
import json, os, sys
args = sys.argv[1:]
if args and os.path.isfile(args[0]):
    text, args = open(args[0]).read(), args[1:]
elif args and args[0] == "-":
    text, args = sys.stdin.read(), args[1:]
else:
    sys.exit("usage: bench_line.py FILE [tag]  |  bench_line.py - [tag] < FILE   (never blocks on a terminal-less stdin by accident)")
lines = [l for l in text.splitlines() if l.startswith("{")]
d = json.loads(lines[-1])
k = d.get("kernels", {})
print(" ".join(args), d["config"]["workload"].split(":")[0], d.get("dtype"), f"N={d['n_gpus']}", d.get("scaling"),
      f"value {d['value']:.4g}", f"ms/step {d['ms_per_step']:.3f}", {n: (round(v["ms"], 3), round(v["frac"], 3)) for n, v in k.items()},
      "launches", d.get("gpu_launches"), "e2e", d.get("e2e", {}).get("value"))
