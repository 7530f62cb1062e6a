#!/usr/bin/env python3
"""Development probe: runs one frame through the CUDA extractor with per-stage synchronisation and prints, per stage,
how far it agrees with the CPU oracle.  Not a test; used on the GPU box while bringing kernels up."""
import os, sys
os.environ.setdefault("ORBX_DEBUG_SYNC", "1")
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np
from oracle import oracle_lib as O
from orb_slam_cuda_b200 import synth
import orb_slam_cuda_b200 as X

kind = sys.argv[1] if len(sys.argv) > 1 else "tum"
nf = {"kitti": 2000, "tum": 1000, "euroc": 1200}[kind]
img = synth.frame(kind, 3)
gx, ox = X.ORBextractor(nf, 1.2, 8, 20, 7), O.Extractor(nf, 1.2, 8, 20, 7)
ok, od = ox.extract(img)
try:
    k, d = gx(img)
except X.OrbxError as e:
    print("EXTRACT FAILED:", e)
    sys.exit(0)
print("n gpu", len(k), "n cpu", len(ok))
t = lambda a: [(float(q["x"]), float(q["y"]), float(q["response"])) for q in a]
for l in range(8):
    a, b = gx.image_pyramid_level(l, padded=True), ox.level_padded(l)
    bl = ox.level_blurred(l)
    gb = gx.debug_blurred_level(l)
    ca, cb = gx.debug_candidates(l), ox.level_candidates(l)
    ka, kb = gx.debug_level_keypoints(l), ox.level_keypoints(l)
    ang = (ka["angle"].view(np.uint32) != kb["angle"].view(np.uint32)).sum() if len(ka) == len(kb) else -1
    print(f"L{l} pyr diff {(a != b).sum()}/{a.size}  blur diff {(gb != bl).sum() if bl is not None else 'n/a'}  "
          f"cand {len(ca)}/{len(cb)} equal={t(ca) == t(cb)}  kept {len(ka)}/{len(kb)} equal={t(ka) == t(kb)} anglediff {ang}")
if len(k) == len(ok):
    for f in ["x", "y", "size", "angle", "response", "octave", "class_id"]:
        print(f, "diff", (k[f].view(np.uint32) != ok[f].view(np.uint32)).sum())
    print("desc bits diff", np.unpackbits(d ^ od).sum(), "of", d.size * 8)
