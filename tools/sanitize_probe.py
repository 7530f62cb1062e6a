#!/usr/bin/env python3
"""compute-sanitizer target (development tool): one pass through every kernel family on small inputs — single-frame and batched
extraction for three shapes (odd sizes included), both pattern variants, stereo matching, the matchers, undistort, all-vs-all.
usage: compute-sanitizer --tool memcheck python tools/sanitize_probe.py"""
import os, sys
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import orb_slam_cuda_b200 as X
from orb_slam_cuda_b200 import synth
rng = np.random.default_rng(1)
for (h, w, nf) in [(376, 1241, 2000), (480, 640, 1000), (97, 131, 200), (61, 64, 100)]:
    ex = X.ORBextractor(nf, 1.2, 8, 20, 7)
    img = synth.textured_frame(5, h, w)
    k, d = ex(img)
    k, d = ex(img)
    k, d = ex(img)  # graph replay
    fr = np.stack([img, np.roll(img, 3, 1), img[::-1].copy()])
    kb, db, cb = ex.extract_batch(fr)
    print(h, w, len(k), cb.tolist())
exL, exR = X.ORBextractor(1200, 1.2, 8, 20, 7), X.ORBextractor(1200, 1.2, 8, 20, 7)
l, r = synth.stereo_pair("euroc", 2)
kl, dl = exL(l); kr, dr = exR(r)
u, z = X.ComputeStereoMatches(exL, exR, kl, dl, kr, dr, 47.9, 435.2)
print("stereo", int((u >= 0).sum()))
m = X.ORBmatcher(0.9, True)
a, b = synth.frame("tum", 1), synth.shifted_frame("tum", 1)
e = X.ORBextractor(1000, 1.2, 8, 20, 7)
(k1, d1), (k2, d2) = e(a), e(b)
k1, d1, k2, d2 = k1.copy(), d1.copy(), k2.copy(), d2.copy()
prev = np.stack([k1["x"], k1["y"]], 1).astype(np.float32)
print("sfi", m.SearchForInitialization(k1, d1, k2, d2, (0.0, 640.0, 0.0, 480.0), prev, 100)[0])
q, dd = synth.descriptor_sets(3, 700, 900)
print("bf", int((m.match_bruteforce(q, dd)[3] >= 0).sum()) if hasattr(m, "match_bruteforce") else "-")
K4, dist = (517.3, 516.5, 318.6, 255.3), (0.2624, -0.9531, -0.0054, 0.0026, 1.1633)
print("und", X.UndistortKeyPoints(k1, K4, dist)["x"][:2])
f2 = X.DeviceFrame(k2, d2, (0.0, 640.0, 0.0, 480.0))
n1 = k1.size
z3 = np.full(n1, 5.0, np.float32)
x3 = np.stack([(k1["x"] + 7 - 320) * z3 / 500, (k1["y"] - 4 - 240) * z3 / 500, z3], 1).astype(np.float32)
sf = (np.float32(1.2) ** np.arange(8, dtype=np.float32)).astype(np.float32)
print("sbp", m.SearchByProjectionFrame(x3, np.ones(n1, np.uint8), np.ones(n1, np.uint8), d1, k1["octave"], k1["angle"], f2, np.zeros(k2.size, np.uint8),
                                       (500.0, 500.0, 320.0, 240.0), 40.0, sf, 15.0, 0)[0])
print("done")
