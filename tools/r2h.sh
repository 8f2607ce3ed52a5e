#!/bin/bash
python -m pytest tests -m gpu -q 2>&1 | tail -80 > gpurun_out/r2h_tests.log
grep -E "passed|failed|FAILED|Error|assert" gpurun_out/r2h_tests.log | head -30
B="--steps 10 --warmup 3 --no-cpu-baseline --no-reference-cuda --no-sustained"
python bench.py $B --workload scaled > gpurun_out/r2h_bench_scaled.json 2> gpurun_out/r2h_bench_scaled.err
python -c "
import json
d=json.load(open('gpurun_out/r2h_bench_scaled.json')); print(round(d['ms_per_step'],4), d['value'], d['kernels_ms_per_step'])" || tail -5 gpurun_out/r2h_bench_scaled.err
