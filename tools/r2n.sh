#!/bin/bash
python -m pytest tests -m gpu -q 2>&1 | tail -40 > gpurun_out/r2n_tests.log
grep -E "passed|failed|FAILED|^E " gpurun_out/r2n_tests.log | head -20
python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-sustained > gpurun_out/r2n_bench.json 2> gpurun_out/r2n_bench.err
python -c "
import json
d=json.load(open('gpurun_out/r2n_bench.json'))
for k in ('ms_per_step','t_full','full_update','full_update_conv_bf16','full_update_graph'): print(k, d.get(k))
print(d['reference_cuda'])" || tail -5 gpurun_out/r2n_bench.err
