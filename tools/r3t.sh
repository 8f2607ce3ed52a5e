#!/bin/bash
for n in 5000 6000 7000 7424; do for tr in 64 128; do
DRM_ROLLOUT_TR=$tr python bench.py --steps 8 --warmup 3 --workload walker --n-states $n --no-cpu-baseline --no-full --no-reference-cuda --no-sustained 2>/dev/null | grep "^{" > gpurun_out/r3t.json
python - $n $tr <<'PY'
import json,sys
d=json.loads(open('gpurun_out/r3t.json').read())
print('n',sys.argv[1],'tile',sys.argv[2], round(d['ms_per_step'],3), round(d['value']/1e6,1))
PY
done; done
