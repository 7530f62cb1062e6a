#!/usr/bin/env python3
"""Per-kernel summary of an `ncu --set full` report (development / evidence tool): one CSV row per profiled launch with
duration, DRAM bytes, issue / pipe utilisation, occupancy and instruction count.
usage: tools/ncu_summary.py report.ncu-rep > profiles/rNN_ncu_full_x.csv
       tools/ncu_summary.py report.ncu-rep --json WORKLOAD FRAMES   also merges per-frame DRAM bytes and warp instructions of every
                                                                  kernel of the captured step into profiles/ncu_latest.json,
                                                                  which bench.py reads for roofline.traffic / roofline.issue"""
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

if "--json" in sys.argv:
    import hashlib, json, os
    i = sys.argv.index("--json")
    workload, frames = sys.argv[i + 1], int(sys.argv[i + 2])
    root = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
    path = os.path.join(root, "profiles", "ncu_latest.json")
    doc = json.load(open(path)) if os.path.exists(path) else {"workloads": {}}
    col = {c: k for k, c in enumerate(h)}
    kern = {}
    for r in rows[2:]:
        name = r[col["Kernel Name"]].split("(")[0].split("::")[-1]
        def num(metric):
            v, u = float(r[col[metric]].replace(",", "")), units[col[metric]]
            return v * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "inst": 1.0, "ms": 1.0, "us": 1e-3, "ns": 1e-6}.get(u, 1.0)
        k = kern.setdefault(name, {"launches": 0, "dram_bytes_per_frame": 0.0, "warp_inst_per_frame": 0.0, "ms": 0.0})
        k["launches"] += 1
        k["dram_bytes_per_frame"] += (num("dram__bytes_read.sum") + num("dram__bytes_write.sum")) / frames
        k["warp_inst_per_frame"] += num("smsp__inst_executed.sum") / frames
        k["ms"] += num("gpu__time_duration.sum")
    src = {f: hashlib.sha256(open(os.path.join(root, "orb_slam_cuda_b200", "csrc", f), "rb").read()).hexdigest()[:12]
           for f in sorted(os.listdir(os.path.join(root, "orb_slam_cuda_b200", "csrc"))) if f.startswith("k_")}
    doc["workloads"][workload] = {"frames_per_launch": frames, "report": os.path.basename(sys.argv[1]), "kernels": kern, "kernel_sources_sha256_12": src,
                                  "what": "ncu --set full --clock-control none of one device-resident step (tools/gpu_run.sh ncu); per-frame = per-launch / frames"}
    json.dump(doc, open(path, "w"), indent=1, sort_keys=True)
