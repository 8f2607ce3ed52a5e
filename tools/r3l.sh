#!/bin/bash
for i in 1 2 3 4 5 6; do
python -m pytest tests/test_optim_gpu.py tests/test_graph_gpu.py -m gpu -q 2>&1 | grep -E "^E  |FAILED|passed|failed" | head -12 | cut -c1-600
done
