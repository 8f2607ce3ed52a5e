#!/bin/bash
python -m pytest tests/test_optim_gpu.py tests/test_graph_gpu.py -m gpu -q 2>&1 | tail -6 | cut -c1-300
python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-sustained 2> gpurun_out/r3k.err | grep "^{" > gpurun_out/r3k_bench_walker.json
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r3k_bench_walker.json').read())
for k in ('full_update','full_update_graph'):
    print(k, {a:b for a,b in (d.get(k) or {}).items() if a in ('ms','replay_sample_ms','conv')})
print(d['reference_cuda'].get('full_update_with_optimizers'))
PY
