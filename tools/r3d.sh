#!/bin/bash
python -m pytest tests/test_observation_gpu.py -m gpu -q -x 2>&1 | tail -15 | cut -c1-300
python -m pytest tests -m gpu -q 2>&1 | tail -40 > gpurun_out/r3d_tests.log
grep -E "passed|failed|FAILED|^E  " gpurun_out/r3d_tests.log | head -30 | cut -c1-300
python bench.py --steps 20 --warmup 5 2> gpurun_out/r3d_walker.err | grep "^{" > gpurun_out/r3d_bench_walker.json
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r3d_bench_walker.json').read())
print(round(d['ms_per_step'],4), round(d['value']/1e6,2), 'e2e', round(d['e2e']['value']/1e6,2))
for k in ('t_full','full_update','full_update_conv_bf16','full_update_graph'):
    print(k, {a:b for a,b in (d.get(k) or {}).items() if a in ('ms','replay_sample_ms','conv')})
print(d['reference_cuda'].get('speedup_vs_best_case'), d['reference_cuda'].get('full_update'), d['reference_cuda'].get('full_update_with_optimizers'))
PY
python tools/full_trace.py > gpurun_out/r3d_trace_full.txt 2>/dev/null; head -14 gpurun_out/r3d_trace_full.txt | cut -c1-150
