"""Instruction mix of the hot kernels from the built library's SASS (cuobjdump -sass): python tools/sass_summary.py > profiles/sass_r02.txt
For each kernel: static instruction count, opcode histogram, and the lines that show what the design leans on (REDUX for the
tournament selection, LDG.E.128 record / row loads, multimem / system-scope stores of the sharded path, MATCH, VOTE)."""
import collections, os, re, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "mvtracker_b200", "libmvt.so")
KERNELS = [("knn_corr_kernel<1, true, 32, true>  (fp32 fused kNN + gather + correlation, deferral build)", "_ZN3mvt15knn_corr_kernelILi1ELb1ELi32ELb1EEEvNS_10CorrParamsE"),
           ("knn_corr_kernel<1, true, 32, false> (one-kernel build)", "_ZN3mvt15knn_corr_kernelILi1ELb1ELi32ELb0EEEvNS_10CorrParamsE"),
           ("knn_corr_kernel<2, true, 32, true>  (bf16 features)", "_ZN3mvt15knn_corr_kernelILi2ELb1ELi32ELb1EEEvNS_10CorrParamsE"),
           ("cloud_build_kernel<float>", "_ZN3mvt18cloud_build_kernelIfEEvNS_11CloudParamsE"),
           ("cloud_build_kernel<__nv_bfloat16>", "_ZN3mvt18cloud_build_kernelI13__nv_bfloat16EEvNS_11CloudParamsE"),
           ("shard_iter_kernel<false> (multi-GPU: producers + tail consumers)", "_ZN3mvt17shard_iter_kernelILb0EEEvNS_11ShardParamsE"),
           ("knn_corr_hard_kernel<1> (tile search for deferred queries)", "_ZN3mvt20knn_corr_hard_kernelILi1EEEvNS_10CorrParamsE")]
WANT = re.compile(r"REDUX|MATCH|LDG\.E\.128|STG\.E\.128|MULTIMEM|\.SYS|ATOM|RED\.|MEMBAR|NANOSLEEP|LDS\.64|STS\.64")
print(f"# cuobjdump -sass {os.path.relpath(LIB, ROOT)}  (sm_100a; summary by tools/sass_summary.py)")
for title, sym in KERNELS:
    out = subprocess.run(["cuobjdump", "-sass", "-fun", sym, LIB], capture_output=True, text=True).stdout
    ops, shown = collections.Counter(), collections.Counter()
    lines = []
    for ln in out.splitlines():
        m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
        if not m:
            continue
        ins = m.group(2).strip()
        ins = re.sub(r"^@!?U?P\d+\s+", "", ins)
        op = ins.split()[0]
        ops[op.split(".")[0]] += 1
        if WANT.search(op) and shown[op] < 2:
            shown[op] += 1
            lines.append(f"    /*{m.group(1)}*/  {ins}")
    n = sum(ops.values())
    print(f"\n== {title}\n   {sym}\n   {n} SASS instructions; " + ", ".join(f"{k} {v}" for k, v in ops.most_common(18)))
    full = collections.Counter()
    for ln in out.splitlines():
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", ln)
        if m and WANT.search(m.group(1)):
            full[m.group(1)] += 1
    print("   notable: " + ", ".join(f"{k} x{v}" for k, v in sorted(full.items())))
    print("\n".join(lines[:24]))
