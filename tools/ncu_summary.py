"""Text summary of an .ncu-rep (raw page) for profiles/: python tools/ncu_summary.py REPORT.ncu-rep [KEY] > profiles/NAME.txt
KEY ("workload/dtype/kernel/nN", e.g. panoptic/f32/knn_corr/n1): also record the capture's DRAM traffic per launch under
that key in profiles/traffic.json -- the table bench.py's roofline.traffic is read from."""
import csv, json, os, subprocess, sys
rep = sys.argv[1]
key = sys.argv[2] if len(sys.argv) > 2 else None
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
h, u = rows[0], rows[1]
want = ["Kernel Name", "Block Size", "Grid Size", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__cycles_active.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct", "lts__t_bytes.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__registers_per_thread", "launch__occupancy_limit_registers", "launch__waves_per_multiprocessor",
        "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "smsp__average_warp_latency_per_inst_issued.ratio", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio", "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio"]
print(f"# ncu --set full --clock-control none, report {rep.split('/')[-1]} (summary by tools/ncu_summary.py)")
for r in rows[2:]:
    for w in want:
        if w in h:
            i = h.index(w)
            print(f"{w:90s} {r[i][:120]} {u[i]}")
    try:
        rd, wr = float(r[h.index('dram__bytes_read.sum')]), float(r[h.index('dram__bytes_write.sum')])
        ur, uw = u[h.index('dram__bytes_read.sum')], u[h.index('dram__bytes_write.sum')]
        sc = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        print(f"{'traffic = dram read + write':90s} {rd * sc[ur] + wr * sc[uw]:.0f} byte")
        if key:
            tj = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "profiles", "traffic.json")
            tab = json.load(open(tj)) if os.path.exists(tj) else {}
            tab[key] = {"bytes": int(rd * sc[ur] + wr * sc[uw]), "capture": rep.split("/")[-1], "kernel": r[h.index("Kernel Name")][:80]}
            json.dump(tab, open(tj, "w"), indent=1, sort_keys=True)
            key = None
    except Exception as e:
        print("traffic: n/a", e)
    print("-" * 60)
