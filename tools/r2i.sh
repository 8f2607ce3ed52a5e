#!/bin/bash
python -m pytest tests -m gpu -q 2>&1 | tail -120 > gpurun_out/r2i_tests.log
grep -E "passed|failed|FAILED|^E " gpurun_out/r2i_tests.log | head -40
python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-sustained > gpurun_out/r2i_bench.json 2> gpurun_out/r2i_bench.err
python -c "
import json
d=json.load(open('gpurun_out/r2i_bench.json'))
for k in ('ms_per_step','t_full','full_update','full_update_conv_bf16','reference_cuda'): print(k, d.get(k))" || tail -5 gpurun_out/r2i_bench.err
