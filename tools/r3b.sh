#!/bin/bash
python tools/step_trace.py --precision fp32 --eager > gpurun_out/r3b_trace_fp32.txt 2>/dev/null
grep -n "wgrad_kernel\|span" gpurun_out/r3b_trace_fp32.txt | head -40 | cut -c1-200
