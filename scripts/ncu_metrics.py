"""Print the handful of ncu metrics the roofline discussion uses, one line per captured launch.
usage: ncu -i X.ncu-rep --page raw --csv | python scripts/ncu_metrics.py"""
import csv
import sys

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tc_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed"]
rows = [r for r in csv.reader(sys.stdin) if r]
hdr = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
names, units = rows[hdr], rows[hdr + 1]
kn = names.index("Kernel Name")
for r in rows[hdr + 2:]:
    if len(r) != len(names):
        continue
    out = [r[kn][:70]]
    for w in WANT:
        if w in names:
            i = names.index(w)
            out.append(f"{w.split('.')[0]}={r[i]} {units[i]}")
    print(" | ".join(out))
