#!/bin/bash
# default build: tests; then the kernel variants one by one on the walker workload (and 8192)
python -m pytest tests -m gpu -x -q 2>&1 | tail -6 > gpurun_out/r2d_tests.log
B="--steps 20 --warmup 5 --no-cpu-baseline --no-reference-cuda --no-full --no-sustained"
for v in _v5 _v6 _v0; do
  for wl in walker; do
    DRM_B200_LIB=dreamer_pytorch_b200/libdreamer_b200$v.so python bench.py $B --workload $wl > gpurun_out/r2d_bench${v}_$wl.json 2> gpurun_out/r2d_bench${v}_$wl.err
  done
done
tail -4 gpurun_out/r2d_tests.log
for f in gpurun_out/r2d_bench*.json; do python -c "
import json,sys
try:
    d=json.load(open('$f')); print('$f', round(d['ms_per_step'],4), d['kernels_ms_per_step'])
except Exception as e: print('$f', 'FAILED')"; done
