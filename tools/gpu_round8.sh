#!/bin/bash
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
python -c "import json; d=json.load(open('gpurun_out/bench.json')); print(json.dumps(d['latency'], indent=1)); print(d['value'], d['e2e']['value'])"; tail -5 gpurun_out/bench.err
