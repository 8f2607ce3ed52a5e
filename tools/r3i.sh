#!/bin/bash
python -m pytest tests/test_peer_allreduce_gpu.py -m gpu -q 2>&1 | tail -8 | cut -c1-400
