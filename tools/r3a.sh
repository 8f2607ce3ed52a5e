#!/bin/bash
python -m pytest tests -m gpu -q -x 2>&1 | tail -40 > gpurun_out/r3a_tests.log
grep -E "passed|failed|FAILED|^E  " gpurun_out/r3a_tests.log | head -30 | cut -c1-300
for spec in "walker fp32" "pong bf16" "scaled bf16" "walker bf16"; do
  set -- $spec
  python bench.py --steps 10 --warmup 3 --workload $1 --precision $2 --no-cpu-baseline --no-sustained 2> gpurun_out/r3a_$1_$2.err | grep "^{" > gpurun_out/r3a_bench_$1_$2.json
  python - "$1" "$2" <<'PY'
import json,sys
try:
    d=json.loads(open(f'gpurun_out/r3a_bench_{sys.argv[1]}_{sys.argv[2]}.json').read())
    print(sys.argv[1], sys.argv[2], round(d['ms_per_step'],3), round(d['value']/1e6,2), 'ref_cuda', (d.get('reference_cuda') or {}).get('speedup_vs_best_case'), 't_full', (d.get('t_full') or {}).get('ms'),
          'full', (d.get('full_update') or {}).get('ms'), 'kern', d.get('kernels_ms_per_step'))
except Exception as e: print(sys.argv[1], 'FAILED', e)
PY
done
