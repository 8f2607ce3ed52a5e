#!/bin/bash
python -m pytest tests -m gpu -q 2>&1 | tail -40 > gpurun_out/r3g_tests.log
grep -E "passed|failed|FAILED|^E  " gpurun_out/r3g_tests.log | head -30 | cut -c1-300
python tools/full_trace.py --rssm-only > gpurun_out/r3g_trace_rssm.txt 2>/dev/null
head -8 gpurun_out/r3g_trace_rssm.txt | cut -c1-150
