#!/bin/bash
python -m pytest tests -m gpu -q 2>&1 | tail -40 > gpurun_out/r2l_tests.log
grep -E "passed|failed|FAILED|^E " gpurun_out/r2l_tests.log | head -20
B="--steps 10 --warmup 3 --no-cpu-baseline --no-reference-cuda --no-sustained"
python bench.py $B > gpurun_out/r2l_bench.json 2> gpurun_out/r2l_bench.err
python bench.py $B --precision fp32 --n-states 512 > gpurun_out/r2l_bench_fp32_512.json 2> gpurun_out/r2l_bench_fp32_512.err
python tools/full_trace.py --rssm-only 2>/dev/null | head -6
for f in gpurun_out/r2l_bench*.json; do python -c "
import json,sys
try:
    d=json.load(open('$f')); print('$f', round(d['ms_per_step'],4), d['kernels_ms_per_step'], d.get('t_full',{}).get('ms'), d.get('full_update',{}).get('ms'))
except Exception as e: print('$f', 'FAILED')"; done
