#!/bin/bash
N=${NG:-8}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 20 --warmup 5 --no-cpu-baseline 2> gpurun_out/r2w_${N}gpu.err | grep "^{" > gpurun_out/r2w_bench_${N}gpu.json
python - $N <<'PY'
import json,sys
n=sys.argv[1]
try:
    d=json.loads(open(f'gpurun_out/r2w_bench_{n}gpu.json').read())
    print(n,'gpus', d['config']['workload'][:50], d['scaling'], round(d['ms_per_step'],3), round(d['value']/1e6,1), 'e2e', round(d['e2e']['value']/1e6,1))
    for k in ('weak','one_gpu_same_workload'):
        print(k, d.get(k))
except Exception as e:
    print('FAILED', e); print(open(f'gpurun_out/r2w_{n}gpu.err').read()[-1500:])
PY
