#!/usr/bin/env python3
"""Development tool: the pieces of bench.py's latency leg with a handful of iterations each and a progress line after every piece,
to be run under ncu when a profiler-only failure has to be located."""
import os, sys, subprocess, tempfile
from concurrent.futures import ThreadPoolExecutor
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import orb_slam_cuda_b200 as X
from orb_slam_cuda_b200 import synth
def mark(s): print("STEP", s, flush=True)
ex = X.ORBextractor(2000, 1.2, 8, 20, 7, 1241, 376, max_batch=1)
imgs = [synth.frame("kitti", 500 + i) for i in range(2)]
for i in range(4): ex(imgs[i % 2])
mark("single frames ok")
with tempfile.NamedTemporaryFile(suffix=".raw") as f:
    f.write(imgs[0].tobytes() + imgs[1].tobytes()); f.flush()
    r = subprocess.run([os.path.join(os.path.dirname(__file__), "..", "tests", "cpp", "frame_harness"), "latency", f.name, "1241", "376", "2000", "3"],
                       capture_output=True, text=True)
mark("harness rc=%d %s" % (r.returncode, r.stdout.strip()[:80]))
left, right = synth.stereo_pair("euroc", 9)
exL, exR = (X.ORBextractor(1200, 1.2, 8, 20, 7, 752, 480, max_batch=1) for _ in range(2))
pool = ThreadPoolExecutor(2); res = {}
def eye(n, e, img): res[n] = e(img)
for i in range(4):
    a, b = pool.submit(eye, "L", exL, left), pool.submit(eye, "R", exR, right); a.result(); b.result()
    u = X.ComputeStereoMatchesLast(exL, exR, len(res["L"][0]), 47.9, 435.2)
mark("stereo ok %d" % int((u[0] >= 0).sum()))
a_, b_ = synth.frame("kitti", 40), synth.shifted_frame("kitti", 40)
(k1, d1), (k2, d2) = ex(a_), ex(b_)
bounds = (0.0, 1241.0, 0.0, 376.0)
m = X.ORBmatcher(0.9, True)
prev = np.stack([k1["x"], k1["y"]], axis=1).astype(np.float32)
for i in range(3): n, m12 = m.SearchForInitialization(k1, d1, k2, d2, bounds, prev.copy(), 100)
mark("sfi ok %d" % n)
pool.shutdown()
mark("done")
