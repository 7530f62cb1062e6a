#!/usr/bin/env python3
"""Single-frame latency probe (development tool): a few orbx_extract calls on one KITTI frame; run under
`ncu --metrics gpu__time_duration.sum` (ORBX_NO_GRAPH=1 so that the kernels are separate launches) for the per-kernel times of ONE frame."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import orb_slam_cuda_b200 as X
from orb_slam_cuda_b200 import synth
kind = sys.argv[1] if len(sys.argv) > 1 else "kitti"
nf = {"kitti": 2000, "tum": 1000, "euroc": 1200}[kind]
h, w = synth.SHAPES[kind]
ex = X.ORBextractor(nf, 1.2, 8, 20, 7, w, h)
img = synth.frame(kind, 5)
ts = []
for i in range(400):
    t0 = time.perf_counter(); ex(img); ts.append((time.perf_counter() - t0) * 1e3)
ts = sorted(ts[20:])
print(kind, "p50 ms", round(ts[len(ts) // 2], 4), "p10", round(ts[len(ts) // 10], 4), "min", round(ts[0], 4))
ex.enable_timing(True)  # events between the stages of a single frame (GPU time of each stage; the call no longer replays a graph)
acc = {}
for i in range(20):
    ex(img)
    if i >= 5:
        for k, v in ex.stage_ms().items():
            acc.setdefault(k, []).append(v)
print(kind, "single-frame stage us:", {k.split()[0]: round(1e3 * sorted(v)[len(v) // 2], 1) for k, v in acc.items()},
      "sum", round(1e3 * sum(sorted(v)[len(v) // 2] for v in acc.values()), 1))
