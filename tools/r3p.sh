#!/bin/bash
python -m pytest tests/test_graph_gpu.py tests/test_noise_gpu.py tests/test_bf16_gpu.py tests/test_optim_gpu.py -m gpu -q 2>&1 | tail -3 | cut -c1-300
python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-full --no-reference-cuda 2> gpurun_out/r3p.err | grep "^{" > gpurun_out/r3p_bench_walker.json
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r3p_bench_walker.json').read())
print(round(d['ms_per_step'],4), round(d['value']/1e6,2), 'e2e', round(d['e2e']['value']/1e6,2), 'sustained', (d.get('sustained') or {}).get('ms_per_step'))
PY
python tools/step_trace.py > gpurun_out/r3p_trace_2500.txt 2>/dev/null; sed -n 34,46p gpurun_out/r3p_trace_2500.txt | cut -c1-130
