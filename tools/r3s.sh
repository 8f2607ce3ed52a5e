#!/bin/bash
python -m pytest tests/test_bf16_gpu.py tests/test_graph_gpu.py -m gpu -q 2>&1 | tail -3 | cut -c1-300
for n in 2500 4096 5000 6000 7000 7424 8192 12288; do for mf in 16 32; do
DRM_B200_MIN_FREE=$mf python bench.py --steps 8 --warmup 3 --workload walker --n-states $n --no-cpu-baseline --no-full --no-reference-cuda --no-sustained 2>/dev/null | grep "^{" > gpurun_out/r3s.json
python - $n $mf <<'PY'
import json,sys
d=json.loads(open('gpurun_out/r3s.json').read())
print('n',sys.argv[1],'minfree',sys.argv[2], round(d['ms_per_step'],3), round(d['value']/1e6,1))
PY
done; done
