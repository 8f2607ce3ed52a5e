#!/bin/bash
python -m pytest tests -m gpu -q 2>&1 | tail -30 > gpurun_out/r2r_tests.log
grep -E "passed|failed|FAILED|^E " gpurun_out/r2r_tests.log | head -20
python tools/full_trace.py --rssm-only > gpurun_out/r2r_trace_rssm.txt 2>/dev/null
head -12 gpurun_out/r2r_trace_rssm.txt | cut -c1-150
python bench.py --steps 20 --warmup 5 2>/dev/null | grep "^{" > gpurun_out/r2r_bench_walker.json
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r2r_bench_walker.json').read())
print({k:d.get(k) for k in ('value','ms_per_step','t_full','full_update','full_update_graph')})
print(d.get('e2e')); print(d.get('reference_cuda'))
PY
