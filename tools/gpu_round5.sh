#!/bin/bash
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
for t in 0 1; do ORBX_PIPE_TAPER=$t python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-match 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('taper $t value', d['value'], 'e2e', d['e2e']['value'], d['stage_ms_per_step'])"; done
python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
cat gpurun_out/bench.json; tail -5 gpurun_out/bench.err
python - <<'PY'
import sys; sys.path.insert(0, '.')
import numpy as np, orb_slam_cuda_b200 as X
from orb_slam_cuda_b200 import synth
ex = X.ORBextractor(2000, 1.2, 8, 20, 7, 1241, 376, max_batch=1)
img = synth.frame("kitti", 3)
ex.enable_timing(True)
for i in range(5): ex(img)
print("single-frame stage ms:", ex.stage_ms())
PY
