#!/bin/bash
# The one GPU-box script (run through gpurun): tools/gpu_run.sh <mode> [args]
#   tests [pytest args]   pytest -m gpu (+ smoke)
#   variants              stage times of every liborbx*.so variant (tools/variants.py builds them)
#   sweep VAR v1 v2 ...   bench the default build with the environment variable VAR set to each value
#   evidence TAG          tests, smoke, both bench arms, tum/euroc lines, ncu launch list, ncu --set full of one step
#   ncu REGEX TAG SKIP N   ncu --set full --import-source of N launches matching REGEX after SKIP of them (device-resident step only)
#   multi N               bench.py under torchrun on N GPUs
mkdir -p gpurun_out
mode=$1; shift
line() {  # one summary line of a bench JSON on stdin
python -c "
import json,sys
d=json.loads(sys.stdin.read()); s=d['stage_ms_per_step']
print('$1', round(d['value']), 'e2e', round(d['e2e']['value']), 'st', d['status'], ' '.join('%.3f'%v for v in s.values()))"
}
case $mode in
tests)
  python -m pytest tests -m gpu -x -q "$@" > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
  tail -15 gpurun_out/pytest_gpu.log
  python -c 'import __graft_entry__ as g; g.smoke()' > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
  ;;
variants)
  for so in orb_slam_cuda_b200/liborbx.so orb_slam_cuda_b200/liborbx_*.so; do
    case "$so" in *classes*) continue;; esac
    ORBX_LIB=$PWD/$so python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-match --no-latency 2>gpurun_out/variant.err | line "$(basename $so)" | tee -a gpurun_out/variants.txt
  done
  ;;
sweep)
  var=$1; shift
  for v in "$@"; do
    env $var=$v python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-match --no-latency 2>gpurun_out/sweep.err | line "$var=$v" | tee -a gpurun_out/sweep.txt
  done
  ;;
evidence)
  tag=${1:-x}
  python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_$tag.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_$tag.log
  tail -4 gpurun_out/pytest_gpu_$tag.log
  python -c 'import __graft_entry__ as g; g.smoke()' > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"
  python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_$tag.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
  python bench.py --steps 10 --warmup 3 > gpurun_out/bench_$tag.json 2> gpurun_out/bench.err; echo "bench rc=$?"
  cat gpurun_out/bench_$tag.json; tail -5 gpurun_out/bench.err
  python bench.py --workload tum --steps 10 --warmup 3 --no-match > gpurun_out/bench_tum_$tag.json 2> gpurun_out/bench_tum.err
  python bench.py --workload euroc --steps 10 --warmup 3 --no-match > gpurun_out/bench_euroc_$tag.json 2> gpurun_out/bench_euroc.err
  python bench.py --steps 2 --warmup 3 --batch 512 --no-cpu-baseline --no-match --no-latency > gpurun_out/plain.log 2>&1 &&
  ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_$tag.csv \
    python bench.py --steps 2 --warmup 3 --batch 512 --no-cpu-baseline --no-match --no-latency > gpurun_out/ncu_launches.log 2>&1
  ;;
ncu)  # ncu KERNEL_REGEX TAG SKIP COUNT [bench args]: --set full with source of the matching launches of one device-resident step
  k=$1; tag=$2; skip=${3:-0}; cnt=${4:-1}; shift 4
  ncu --set full --import-source on --clock-control none -k regex:$k -s $skip -c $cnt -f -o gpurun_out/ncu_$tag \
    python bench.py --profile --steps 1 --warmup 1 --batch 512 "$@" > gpurun_out/ncu_$tag.log 2>&1; echo "ncu rc=$?"
  ls -la gpurun_out/ncu_$tag.ncu-rep
  ;;
ncusum)  # ncusum WORKLOAD TAG [keep]: capture one step of WORKLOAD, write the per-kernel csv + profiles/ncu_latest.json entry, drop the
         # 20+ MB report unless `keep` is given (gpurun brings back at most 64 MiB)
  wl=$1; tag=$2; keep=$3
  ncu --set full --import-source on --clock-control none -k regex:k_ -s 13 -c 13 -f -o gpurun_out/ncu_${tag}_$wl \
    python bench.py --profile --steps 1 --warmup 1 --batch 512 --workload $wl > gpurun_out/ncu_${tag}_$wl.log 2>&1; echo "ncu $wl rc=$?"
  python tools/ncu_summary.py gpurun_out/ncu_${tag}_$wl.ncu-rep --json $wl 512 > gpurun_out/ncu_full_${tag}_$wl.csv
  cp profiles/ncu_latest.json gpurun_out/ncu_latest.json
  [ "$keep" = keep ] || rm -f gpurun_out/ncu_${tag}_$wl.ncu-rep
  ;;
multi)
  n=$1
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29512 tools/multi_gpu_check.py \
    > gpurun_out/multi_check_n$n.json 2> gpurun_out/multi_check_n$n.err; echo "check rc=$?"; cat gpurun_out/multi_check_n$n.json; tail -3 gpurun_out/multi_check_n$n.err
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --steps 10 --warmup 3 \
    > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err; echo "rc=$?"; cat gpurun_out/bench_n$n.json; tail -3 gpurun_out/bench_n$n.err
  ;;
esac
