#!/bin/bash
# N-GPU check of the sharded workloads: bench.py under torchrun, both arms
N=${1:-2}
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -k "graph_replay or chunked" 2>&1 | tail -3
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "rc=$?"
cat gpurun_out/bench_n$N.json; tail -5 gpurun_out/bench_n$N.err
python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n1.json 2>/dev/null; cat gpurun_out/bench_n1.json
