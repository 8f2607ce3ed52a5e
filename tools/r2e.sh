#!/bin/bash
python -m pytest tests -m gpu -x -q 2>&1 | tail -6 > gpurun_out/r2e_tests.log
B="--steps 20 --warmup 5 --no-cpu-baseline --no-reference-cuda --no-sustained"
python bench.py $B > gpurun_out/r2e_bench.json 2> gpurun_out/r2e_bench.err
DRM_B200_LIB=dreamer_pytorch_b200/libdreamer_b200_vA.so python bench.py $B --no-full > gpurun_out/r2e_bench_vA.json 2> gpurun_out/r2e_bench_vA.err
python tools/full_trace.py > gpurun_out/r2e_trace_full.txt 2>/dev/null
tail -4 gpurun_out/r2e_tests.log
for f in gpurun_out/r2e_bench*.json; do python -c "
import json,sys
try:
    d=json.load(open('$f')); print('$f', round(d['ms_per_step'],4), d['kernels_ms_per_step'], d.get('t_full'), d.get('full_update'))
except Exception as e: print('$f', 'FAILED')"; done
head -20 gpurun_out/r2e_trace_full.txt
