#!/bin/bash
for tr in 64 128; do for n in 16384 20000; do
DRM_ROLLOUT_TR=$tr python bench.py --steps 10 --warmup 3 --workload walker --n-states $n --no-cpu-baseline --no-full --no-reference-cuda --no-sustained 2>/dev/null | grep "^{" > gpurun_out/r3q_${n}_${tr}.json
python - $n $tr <<'PY'
import json,sys
d=json.loads(open(f'gpurun_out/r3q_{sys.argv[1]}_{sys.argv[2]}.json').read())
print('n',sys.argv[1],'tile',sys.argv[2], round(d['ms_per_step'],3), round(d['value']/1e6,1))
PY
done; done
