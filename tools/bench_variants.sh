#!/bin/bash
# Development aid (GPU box): stage times of every liborbx_<name>.so variant next to the default build.
for so in orb_slam_cuda_b200/liborbx.so orb_slam_cuda_b200/liborbx_*.so; do
  case "$so" in *classes*) continue;; esac
  ORBX_LIB=$PWD/$so python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-match 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); s=d['stage_ms_per_step']
print('$so'.split('/')[-1], round(d['value']), 'st', d['status'], ' '.join('%.3f'%v for v in s.values()))"
done
