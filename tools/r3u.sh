#!/bin/bash
python -m pytest tests -m gpu -q 2>&1 | tail -40 > gpurun_out/r3u_tests.log
grep -E "passed|failed|FAILED|^E  " gpurun_out/r3u_tests.log | head -20 | cut -c1-300
for n in 2500 5000 6000 8192 16384; do
python bench.py --steps 8 --warmup 3 --workload walker --n-states $n --no-cpu-baseline --no-full --no-reference-cuda --no-sustained 2>/dev/null | grep "^{" > gpurun_out/r3u.json
python - $n <<'PY'
import json,sys
d=json.loads(open('gpurun_out/r3u.json').read())
print('n',sys.argv[1], round(d['ms_per_step'],3), round(d['value']/1e6,1))
PY
done
