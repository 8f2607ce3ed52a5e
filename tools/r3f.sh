#!/bin/bash
python -m pytest tests/test_ops_gpu.py -m gpu -q -x 2>&1 | tail -25 | cut -c1-300
