#!/usr/bin/env python3
"""Warp instructions per SASS address range of an `ncu --page source --csv --print-source cuda,sass` dump (development tool).
usage: ncu_phases.py src.csv name:lo-hi ...   (hex offsets from the kernel's first instruction)"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = next(r for r in rows if r and r[0] == "Line No")
ai, ii = hdr.index("Address"), hdr.index("Instructions Executed")
si = hdr.index("Warp Stall Sampling (All Samples)")
sass = [(int(r[ai], 16), int(r[ii]), int(r[si]) if r[si].isdigit() else 0, r[3]) for r in rows
        if len(r) > ii and r[0] == "" and r[ai].startswith("0x") and r[ii].isdigit()]
seen = {}
for a, n, s, t in sass:
    seen[a] = (n, s, t)
base = min(seen)
tot = sum(v[0] for v in seen.values()); tots = sum(v[1] for v in seen.values())
print("total", tot, "stall samples", tots, "instructions", len(seen))
for spec in sys.argv[2:]:
    name, rng = spec.split(":")
    lo, hi = (int(x, 16) for x in rng.split("-"))
    n = sum(v[0] for a, v in seen.items() if lo <= a - base < hi)
    s = sum(v[1] for a, v in seen.items() if lo <= a - base < hi)
    print(f"{name:10s} inst {n / tot * 100:5.1f}%  stall {s / max(tots,1) * 100:5.1f}%  ({n})")
