#!/bin/bash
python -m pytest tests -m gpu -x -q 2>&1 | tail -6 > gpurun_out/r2f_tests.log
B="--steps 10 --warmup 3 --no-cpu-baseline --no-reference-cuda --no-sustained"
for v in "" _k0 _k8; do
  DRM_B200_LIB=dreamer_pytorch_b200/libdreamer_b200$v.so python bench.py $B > gpurun_out/r2f_bench$v.json 2> gpurun_out/r2f_bench$v.err
  DRM_B200_LIB=dreamer_pytorch_b200/libdreamer_b200$v.so python bench.py $B --precision fp32 > gpurun_out/r2f_bench${v}_fp32.json 2> gpurun_out/r2f_bench${v}_fp32.err
done
tail -4 gpurun_out/r2f_tests.log
for f in gpurun_out/r2f_bench*.json; do python -c "
import json,sys
try:
    d=json.load(open('$f')); print('$f', round(d['ms_per_step'],4), d['kernels_ms_per_step'], d.get('t_full',{}).get('ms'), d.get('full_update',{}).get('ms'))
except Exception as e: print('$f', 'FAILED')"; done
