#!/bin/bash
for n in 6000 8192 10240; do for sch in "heads,prepare,backward" "prepare" "plain"; do
DRM_B200_SCHEDULE=$sch python bench.py --steps 8 --warmup 3 --workload walker --n-states $n --no-cpu-baseline --no-full --no-reference-cuda --no-sustained 2>/dev/null | grep "^{" > gpurun_out/r3v.json
python - $n "$sch" <<'PY'
import json,sys
d=json.loads(open('gpurun_out/r3v.json').read())
print('n',sys.argv[1],'schedule',sys.argv[2], round(d['ms_per_step'],3), round(d['value']/1e6,1))
PY
done; done
