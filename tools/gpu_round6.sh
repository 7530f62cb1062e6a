#!/bin/bash
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -8 gpurun_out/pytest_gpu.log
python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
cat gpurun_out/bench.json; tail -5 gpurun_out/bench.err
ncu --set full --clock-control none --import-source on -s 36 -c 12 -o gpurun_out/round6_full -f \
  python bench.py --steps 1 --warmup 3 --batch 512 --no-cpu-baseline --no-match > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out
