"""Per-source-line instruction counts of one kernel in an .ncu-rep (needs the cubin the report was taken from).
    python tools/ncu_lines.py REPORT.ncu-rep LIB.so KERNEL_SUBSTR [N_UNITS]
"""
import collections, csv, re, subprocess, sys, tempfile, os
rep, so, kern = sys.argv[1:4]
units = float(sys.argv[4]) if len(sys.argv) > 4 else 1.0
tmp = tempfile.mkdtemp()
subprocess.check_call(["cuobjdump", "-xelf", "all", os.path.abspath(so)], cwd=tmp, stdout=subprocess.DEVNULL)
addr2line = {}
for f in os.listdir(tmp):
    if not f.endswith(".cubin") or f.count("-") > 0:
        continue
    dis = subprocess.run(["nvdisasm", "--print-line-info", os.path.join(tmp, f)], capture_output=True, text=True).stdout
    cur = line = None
    for l in dis.splitlines():
        m = re.match(r'\s*\.section\s+\.text\.(\S+),', l)
        if m: cur = m.group(1); continue
        m = re.match(r'\s*//## File "([^"]+)", line (\d+)', l)
        if m: line = (m.group(1).split('/')[-1], int(m.group(2))); continue
        m = re.match(r'\s*/\*([0-9a-f]{4,})\*/\s+(.*?);', l)
        if m and cur and kern in cur:
            addr2line[(cur, int(m.group(1), 16))] = (line, m.group(2))
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
starts = [i for i, r in enumerate(rows) if r and r[0] == 'Kernel Name']
hdr = rows[starts[0] + 1]
iAd, iI, iS, iSrc = hdr.index('Address'), hdr.index('Instructions Executed'), hdr.index('# Samples'), hdr.index('Source')
body = rows[starts[0] + 2:(starts[1] if len(starts) > 1 else len(rows))]
base = int(body[0][iAd], 16)
fn = [c for (c, a) in addr2line if a == 0]
fn = fn[0] if len(set(fn)) == 1 else None
byline, samp, byop = collections.Counter(), collections.Counter(), collections.Counter()
for r in body:
    try: n, s = int(r[iI]), int(r[iS])
    except Exception: continue
    a = int(r[iAd], 16) - base
    cands = [v for (c, aa), v in addr2line.items() if aa == a] if fn is None else [addr2line.get((fn, a))]
    ln = cands[0][0] if cands and cands[0] else None
    byline[ln] += n; samp[ln] += s
    op = r[iSrc].split()[1] if r[iSrc].startswith('@') else r[iSrc].split()[0]
    byop[op.split('.')[0]] += n
tot, ts = sum(byline.values()), max(sum(samp.values()), 1)
print(f"total warp instructions {tot}  ({tot / units:.1f} per unit)")
print("top opcodes:", ", ".join(f"{o} {n / units:.0f}" for o, n in byop.most_common(14)))
byfile = collections.Counter()
for k, n in byline.items(): byfile[k[0] if k else None] += n
print("by file:", ", ".join(f"{f} {n / units:.0f}" for f, n in byfile.most_common()))
for k, n in (sorted(byline.items(), key=lambda kv: (str(kv[0]))) if os.environ.get("BYLINE") else byline.most_common(45)):
    print(f"{str(k):38s} {n / units:8.1f}/unit {100 * n / tot:5.1f}%   stall samples {100 * samp[k] / ts:5.1f}%")
