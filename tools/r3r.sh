#!/bin/bash
for n in 8192 10240 12288 14336 16384; do for tr in 64 128; do for mf in 16 40; do
DRM_B200_MIN_FREE=$mf DRM_ROLLOUT_TR=$tr python bench.py --steps 8 --warmup 3 --workload walker --n-states $n --no-cpu-baseline --no-full --no-reference-cuda --no-sustained 2>/dev/null | grep "^{" > gpurun_out/r3r.json
python - $n $tr $mf <<'PY'
import json,sys
d=json.loads(open('gpurun_out/r3r.json').read())
print('n',sys.argv[1],'tile',sys.argv[2],'minfree',sys.argv[3], round(d['ms_per_step'],3), round(d['value']/1e6,1))
PY
done; done; done
