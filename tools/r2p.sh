#!/bin/bash
python -m pytest tests -m gpu -q 2>&1 | tail -30 > gpurun_out/r2p_tests.log
grep -E "passed|failed|FAILED|^E " gpurun_out/r2p_tests.log | head -20
python tools/step_trace.py --n 8192 > gpurun_out/r2p_trace_8192.txt 2>/dev/null
cat gpurun_out/r2p_trace_8192.txt | cut -c1-150
