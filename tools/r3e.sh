#!/bin/bash
cat > /tmp/bm.py <<'PY'
import os, sys, runpy, torch
torch.backends.cudnn.benchmark = True
sys.argv = ["bench.py", "--steps", "10", "--warmup", "3", "--no-cpu-baseline", "--no-sustained"]
runpy.run_path("bench.py", run_name="__main__")
PY
python /tmp/bm.py 2> gpurun_out/r3e.err | grep "^{" > gpurun_out/r3e_bench_cudnn_benchmark.json
python - <<'PY'
import json
d=json.loads(open('gpurun_out/r3e_bench_cudnn_benchmark.json').read())
for k in ('t_full','full_update','full_update_conv_bf16','full_update_graph'):
    print(k, {a:b for a,b in (d.get(k) or {}).items() if a in ('ms','replay_sample_ms','conv')})
print(d['reference_cuda'].get('full_update'), d['reference_cuda'].get('full_update_with_optimizers'))
PY
tail -3 gpurun_out/r3e.err
