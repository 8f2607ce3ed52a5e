#!/bin/bash
bash tools/r2z.sh
python bench.py 2> gpurun_out/r3w.err | grep "^{" > gpurun_out/r3w_bench_walker.json
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r3w_bench_walker.json').read())
print(round(d['ms_per_step'],4), round(d['value']/1e6,2), 'e2e', round(d['e2e']['value']/1e6,2), 'launches', d.get('gpu_launches'), 'roof', d['roofline']['frac'], d['clocks'])
for k in ('t_full','full_update','full_update_graph'):
    print(k, {a:b for a,b in (d.get(k) or {}).items() if a in ('ms','replay_sample_ms','conv')})
print(d['reference_cuda'].get('speedup_vs_best_case'), d['cpu_baseline']['value'], d['cpu_baseline']['kind'])
PY
