#!/usr/bin/env python
"""Print the handful of ncu metrics that matter for the stage kernel from a .ncu-rep file."""
import csv, io, subprocess, sys
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "launch__registers_per_thread", "launch__grid_size", "sm__warps_active.avg.per_cycle_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "sm__cycles_elapsed.max", "lts__t_sector_hit_rate.pct",
        "l1tex__t_sector_hit_rate.pct", "smsp__warps_eligible.avg.per_cycle_active"]
STALL = "smsp__average_warps_issue_stalled_"
for r in rows[2:]:
    print(r[hdr.index("Kernel Name")][:80])
    for k in KEYS:
        if k in hdr:
            print(f"   {k:70s} {r[hdr.index(k)]} {units[hdr.index(k)]}")
    st = [(float(r[i]), k[len(STALL):-len('_per_issue_active.ratio')]) for i, k in enumerate(hdr) if k.startswith(STALL) and k.endswith("_per_issue_active.ratio")]
    print("   stalls per issue:", ", ".join(f"{n}={v:.2f}" for v, n in sorted(st, reverse=True)[:7]))
