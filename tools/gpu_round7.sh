#!/bin/bash
# evidence round: tests, smoke, both bench arms, ncu launch list, ncu --set full of one 512-frame step
set -x
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
python -c 'import __graft_entry__ as g; g.smoke()' > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
cat gpurun_out/bench_ref.json
python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
cat gpurun_out/bench.json; tail -5 gpurun_out/bench.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
  python bench.py --steps 2 --warmup 3 --batch 512 --no-cpu-baseline --no-match > gpurun_out/ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -s 36 -c 12 -o gpurun_out/round7_full -f \
  python bench.py --steps 1 --warmup 3 --batch 512 --no-cpu-baseline --no-match > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out | tail -8
