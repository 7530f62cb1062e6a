#!/usr/bin/env python3
"""Writes tests/golden/ref_matcher.npz: the outputs of the REFERENCE's own ORBmatcher / Frame object code
(oracle/_ref/libref_matcher.so, compiled from /root/reference by oracle/ref_matcher/Makefile) on the seeded cases of
tests/ref_matcher_cases.py.  Run in a container that has /root/reference; the fixture lets every other box check the oracle
against the reference without it (tests/test_oracle_ref_matcher.py::test_oracle_against_reference_golden)."""
import os
import sys

import numpy as np

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import ref_matcher_cases as RC  # noqa: E402

assert RC.R.available(), "the reference library could not be built (no /root/reference?)"
out = {}
for name, make in RC.CASES.items():
    _, reference = make()
    for i, a in enumerate(reference()):
        out[f"{name}/{i}"] = a
    print(name, [out[k].shape for k in out if k.startswith(name + "/")])
path = os.path.join(ROOT, "tests", "golden", "ref_matcher.npz")
np.savez_compressed(path, **out)
print(path, os.path.getsize(path), "bytes")
