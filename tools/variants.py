#!/usr/bin/env python3
"""Development aid: build kernel variants (extra -D flags) and print the bench command that compares them on the GPU box.
usage: tools/variants.py name1:DEF=V,DEF2=V name2:...   (variants are gitignored .so files that travel with gpurun)"""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from orb_slam_cuda_b200 import build as B
from concurrent.futures import ThreadPoolExecutor
specs = [a.split(":", 1) for a in sys.argv[1:]]
with ThreadPoolExecutor(8) as ex:
    outs = list(ex.map(lambda s: B.build_variant(s[0], [d for d in s[1].split(",") if d]), specs))
for o in outs:
    print(o)
