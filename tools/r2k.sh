#!/bin/bash
set -o pipefail
CMD="python bench.py --steps 2 --warmup 1 --no-graph --no-cpu-baseline --no-reference-cuda --no-full --no-sustained --no-secondary"
$CMD > gpurun_out/r2_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2_bf16_launches.csv $CMD > gpurun_out/r2_ncu1.log 2>&1
python tools/prof_r2.py 2 > gpurun_out/r2_plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:rollout_.*kernel -s 4 -c 4 -o gpurun_out/r2_rollouts python tools/prof_r2.py 2 > gpurun_out/r2_ncu2.log 2>&1
tail -3 gpurun_out/r2_ncu1.log gpurun_out/r2_ncu2.log; ls -la gpurun_out/*.ncu-rep gpurun_out/r2_bf16_launches.csv
