#!/bin/bash
# GPU-box recipe (iteration): parity tests, bench, variants, ncu of FAST + describe
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
cat gpurun_out/bench.json; tail -5 gpurun_out/bench.err
bash tools/bench_variants.sh > gpurun_out/variants.txt 2>&1; cat gpurun_out/variants.txt
for c in 32 128; do ORBX_PIPE_CHUNK=$c python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-match 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('chunk $c e2e', d['e2e'])"; done
ncu --set full --clock-control none --import-source on -k regex:'k_fast_strips|k_orient_describe|k_quadtree' -s 6 -c 3 -o gpurun_out/round2_full -f \
  python bench.py --steps 1 --warmup 3 --batch 512 --no-cpu-baseline --no-match > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out
