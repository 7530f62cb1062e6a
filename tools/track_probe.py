"""Runs the tracking matchers of bench.py's latency leg a few times (for `ncu --metrics gpu__time_duration.sum`)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import orb_slam_cuda_b200 as X
from orb_slam_cuda_b200 import synth
import bench

a, b = synth.frame("kitti", 40), synth.shifted_frame("kitti", 40)
ex2 = X.ORBextractor(2000, 1.2, 8, 20, 7, 1241, 376, max_batch=1)
(k1, d1), (k2, d2) = ex2(a), ex2(b)
k1, d1, k2, d2 = k1.copy(), d1.copy(), k2.copy(), d2.copy()


def p50(fn, n=int(os.environ.get("PROBE_N", "30")), warm=3):
    for _ in range(warm):
        fn()
    t = []
    for _ in range(n):
        t0 = time.perf_counter(); fn(); t.append((time.perf_counter() - t0) * 1e3)
    return float(np.sort(np.array(t))[n // 2])


out = bench.tracking_leg(X, k1, d1, k2, d2, (0.0, 1241.0, 0.0, 376.0), p50, False)
for k, v in out.items():
    print(k, v["p50_ms"])
