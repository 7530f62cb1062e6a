#!/usr/bin/env python3
"""Per-kernel summary of an `ncu --set full` report (development / evidence tool): one CSV row per profiled launch with
duration, DRAM bytes, issue / pipe utilisation, occupancy and instruction count.
usage: tools/ncu_summary.py report.ncu-rep > profiles/rNN_ncu_full_x.csv"""
import csv, subprocess, sys
M = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
     "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
     "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
     "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
     "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic", "launch__shared_mem_per_block_static",
     "smsp__inst_executed.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"]
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv", "--metrics", ",".join(M)], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
h, units = rows[0], rows[1]
keep = [i for i, c in enumerate(h) if c in ("Kernel Name", "Grid Size", "Block Size") or c in M]
w = csv.writer(sys.stdout)
w.writerow([h[i] + (f" [{units[i]}]" if units[i] else "") for i in keep])
for r in rows[2:]:
    w.writerow([r[i].split("(")[0] if h[i] == "Kernel Name" else r[i] for i in keep])
