#!/bin/bash
B="--steps 5 --warmup 3 --no-cpu-baseline --no-reference-cuda --no-sustained --no-full --precision fp32"
for cap in 0 8; do for wl in walker walker_large pong; do
DRM_FFMA_MAX_ROWS=$cap python bench.py $B --workload $wl > gpurun_out/r2m_${wl}_$cap.json 2> gpurun_out/r2m_${wl}_$cap.err
python -c "
import json,sys
try:
    d=json.load(open('gpurun_out/r2m_${wl}_$cap.json')); print('$wl', $cap, round(d['ms_per_step'],4), d['kernels_ms_per_step'])
except Exception as e: print('$wl $cap FAILED')"
done; done
