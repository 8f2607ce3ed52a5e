#!/bin/bash
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 10 --warmup 3 --workload walker > gpurun_out/r2j_bench_2gpu_weak.json 2> gpurun_out/r2j_bench_2gpu_weak.err
grep "^{" gpurun_out/r2j_bench_2gpu_weak.json | python -c "
import json,sys
d=json.loads(sys.stdin.read())
for k in ('value','ms_per_step','scaling','sustained'): print(k, d.get(k))
print(d['config']['parallelism'])" || tail -30 gpurun_out/r2j_bench_2gpu_weak.err
timeout 120 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --impl reference --gpus 2 --steps 3 --warmup 3 | grep "^{" | cut -c1-400
