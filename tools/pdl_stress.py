#!/usr/bin/env python3
"""Stress check of the single-frame launch sequence (graph replay + programmatic dependent launch + single-frame kernel shapes):
N distinct frames are extracted once as one batch (the throughput shapes, plain launches) and then `passes` times one by one
through orbx_extract, alternating two extractors on two host threads as the stereo path does; every single-frame result must
equal the batch result byte for byte.  Development tool: python tools/pdl_stress.py [workload] [frames] [passes]"""
import os, sys, threading
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import orb_slam_cuda_b200 as X
from orb_slam_cuda_b200 import synth
kind = sys.argv[1] if len(sys.argv) > 1 else "kitti"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 48
passes = int(sys.argv[3]) if len(sys.argv) > 3 else 20
nf = {"kitti": 2000, "tum": 1000, "euroc": 1200}[kind]
h, w = synth.SHAPES[kind]
frames = synth.frame_batch(kind, n, pool=n, first=300)
ref = X.ORBextractor(nf, 1.2, 8, 20, 7, w, h, max_batch=n)
kps, desc, counts = ref.extract_batch(frames)
bad = [0, 0]
def worker(t):
    ex = X.ORBextractor(nf, 1.2, 8, 20, 7, w, h)
    for p in range(passes):
        for f in range(t, n, 2):
            k, d = ex(frames[f])
            c = counts[f]
            if len(k) != c or not np.array_equal(k.view(np.uint8), kps[f, :c].view(np.uint8)) or not np.array_equal(d, desc[f, :c]):
                bad[t] += 1
th = [threading.Thread(target=worker, args=(t,)) for t in range(2)]
[t.start() for t in th]; [t.join() for t in th]
print(kind, "single-frame calls:", n * passes, "mismatches:", sum(bad))
sys.exit(1 if sum(bad) else 0)
