#!/bin/bash
python -m pytest tests -m gpu -q 2>&1 | tail -60 > gpurun_out/r2g_tests.log
B="--steps 10 --warmup 3 --no-cpu-baseline --no-reference-cuda --no-sustained"
python bench.py $B > gpurun_out/r2g_bench.json 2> gpurun_out/r2g_bench.err
python bench.py $B --precision fp32 > gpurun_out/r2g_bench_fp32.json 2> gpurun_out/r2g_bench_fp32.err
python bench.py $B --workload pong > gpurun_out/r2g_bench_pong.json 2> gpurun_out/r2g_bench_pong.err
python bench.py $B --workload scaled > gpurun_out/r2g_bench_scaled.json 2> gpurun_out/r2g_bench_scaled.err
python bench.py $B --workload scaled --precision fp32 > gpurun_out/r2g_bench_scaled_fp32.json 2> gpurun_out/r2g_bench_scaled_fp32.err
python tools/full_trace.py --rssm-only > gpurun_out/r2g_trace_rssm.txt 2>/dev/null
grep -E "passed|failed|FAILED|Error|assert" gpurun_out/r2g_tests.log | head -20
for f in gpurun_out/r2g_bench*.json; do python -c "
import json,sys
try:
    d=json.load(open('$f')); print('$f', round(d['ms_per_step'],4), d['kernels_ms_per_step'], d.get('t_full',{}).get('ms'), d.get('full_update',{}).get('ms'))
except Exception as e: print('$f', 'FAILED'); print(open('$f'.replace('.json','.err')).read()[-600:])"; done
head -8 gpurun_out/r2g_trace_rssm.txt
