#!/bin/bash
python -m pytest tests/test_peer_allreduce_gpu.py -m gpu -q 2>&1 | tail -12 | cut -c1-400
NG=4 bash tools/r2w.sh
grep -o '"parallelism": "[^"]*"' gpurun_out/r2w_bench_4gpu.json
