#!/usr/bin/env python3
"""Writes tests/golden/ref_extractor.npz: keypoints and descriptors produced by the REFERENCE's own ORBextractor object code
(oracle/_ref/libref_extractor.so, compiled from /root/reference by oracle/ref_extractor/Makefile) on the seeded frames of
tests/ref_extractor_cases.py.  Run in a container that has /root/reference; the fixture lets every other box (the GPU box
included) check the oracle and the CUDA extractor against the reference without it."""
import os
import sys

import numpy as np

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import ref_extractor_cases as RC  # noqa: E402
from oracle.ref_extractor import ref_extractor as R  # noqa: E402

assert R.available(), "the reference library could not be built (no /root/reference?)"
out = {}
for name in RC.GOLDEN:
    make, kw = RC.CASES[name]
    img = make()
    h, w = img.shape
    k, d = R.RefExtractor(w, h, **kw).extract(img)
    k6 = np.stack([k["x"], k["y"], k["size"], k["angle"], k["response"], k["octave"].astype(np.float32)], axis=1).astype(np.float32)
    out[name + "/kps"] = k6
    out[name + "/desc"] = d
    out[name + "/img_sum"] = np.int64(img.astype(np.int64).sum())
    print(name, k6.shape, d.shape)
path = os.path.join(ROOT, "tests", "golden", "ref_extractor.npz")
np.savez_compressed(path, **out)
print(path, os.path.getsize(path), "bytes")
