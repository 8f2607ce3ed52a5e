#!/bin/bash
# tests + bench of the batch-4 (default) and batch-2 builds
python -m pytest tests -m gpu -x -q 2>&1 | tail -6 > gpurun_out/r2b_tests.log
B="--steps 20 --warmup 5 --no-cpu-baseline --no-reference-cuda --no-full"
python bench.py $B > gpurun_out/r2b_bench_b4.json 2>/dev/null
DRM_B200_LIB=dreamer_pytorch_b200/libdreamer_b200_b2.so python bench.py $B > gpurun_out/r2b_bench_b2.json 2>/dev/null
python bench.py $B --workload walker_large > gpurun_out/r2b_bench_b4_8192.json 2>/dev/null
DRM_B200_LIB=dreamer_pytorch_b200/libdreamer_b200_b2.so python bench.py $B --workload walker_large > gpurun_out/r2b_bench_b2_8192.json 2>/dev/null
python bench.py $B --workload config4 > gpurun_out/r2b_bench_b4_65536.json 2>/dev/null
tail -4 gpurun_out/r2b_tests.log
for f in gpurun_out/r2b_bench_*.json; do echo $f; python -c "import json,sys; d=json.load(open('$f')); print(d['ms_per_step'], d['value'], d['kernels_ms_per_step'])"; done
