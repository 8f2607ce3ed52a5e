#!/bin/bash
python -m pytest tests/test_parity_gpu.py tests/test_api_gpu.py tests/test_noise_gpu.py -m gpu -q 2>&1 | tail -3 | cut -c1-300
for spec in "scaled bf16" "walker fp32"; do
  set -- $spec
  python bench.py --steps 10 --warmup 3 --workload $1 --precision $2 --no-cpu-baseline --no-sustained --no-full 2> gpurun_out/r3n_$1_$2.err | grep "^{" > gpurun_out/r3n_bench_$1_$2.json
  python - "$1" "$2" <<'PY'
import json,sys
try:
    d=json.loads(open(f'gpurun_out/r3n_bench_{sys.argv[1]}_{sys.argv[2]}.json').read())
    print(sys.argv[1], sys.argv[2], round(d['ms_per_step'],3), round(d['value']/1e6,2), 'ref_cuda', (d.get('reference_cuda') or {}).get('speedup_vs_best_case'), 'kern', d.get('kernels_ms_per_step'))
except Exception as e: print(sys.argv[1], 'FAILED', e)
PY
done
