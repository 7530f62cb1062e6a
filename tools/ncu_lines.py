#!/usr/bin/env python3
"""Per-source-line instruction shares from `ncu --page source --csv --print-source cuda,sass` (development tool)."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
thr = float(sys.argv[2]) if len(sys.argv) > 2 else 0.01
hdr = next(r for r in rows if r and r[0] == "Line No")
ii, ti, wi = hdr.index("Instructions Executed"), hdr.index("Avg. Threads Executed"), hdr.index("Warp Stall Sampling (All Samples)")
lines = [r for r in rows if r and r[0].isdigit() and len(r) > ii and r[ii].isdigit()]
tot = sum(int(r[ii]) for r in lines)
tots = sum(int(r[wi]) for r in lines if r[wi].isdigit())
print("total warp instructions", tot, "stall samples", tots)
for r in lines:
    n = int(r[ii])
    s = int(r[wi]) if r[wi].isdigit() else 0
    if n > tot * thr or s > tots * thr:
        print(f"{r[0]:>4} inst {n / tot * 100:5.1f}%  stall {s / max(tots, 1) * 100:5.1f}%  thr {r[ti]:>3}  {r[1].strip()[:100]}")
