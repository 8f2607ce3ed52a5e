#!/bin/bash
B="--steps 10 --warmup 3 --no-cpu-baseline --no-reference-cuda --no-sustained --no-full"
for tr in 64 128; do for n in 4096 8192 12288; do
DRM_ROLLOUT_TR=$tr python bench.py $B --workload walker --n-states $n > gpurun_out/r2o_${n}_$tr.json 2> gpurun_out/r2o_${n}_$tr.err
python -c "
import json,sys
try:
    d=json.load(open('gpurun_out/r2o_${n}_$tr.json')); print($n, $tr, round(d['ms_per_step'],4), round(d['value']/1e6,1), d['kernels_ms_per_step'])
except Exception as e: print('$n $tr FAILED')"
done; done
