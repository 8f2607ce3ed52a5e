#!/bin/bash
python -m pytest tests/test_optim_gpu.py tests/test_graph_gpu.py -m gpu -q 2>&1 | tail -3 | cut -c1-300
bash tools/r3j.sh
