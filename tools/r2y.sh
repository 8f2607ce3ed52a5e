#!/bin/bash
python -m pytest tests/test_peer_allreduce_gpu.py -m gpu -q -x 2>&1 | tail -30 | cut -c1-400
NG=2 bash tools/r2w.sh
grep -o '"parallelism": "[^"]*"' gpurun_out/r2w_bench_2gpu.json
tail -5 gpurun_out/r2w_2gpu.err | cut -c1-300
