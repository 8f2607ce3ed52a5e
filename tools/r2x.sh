#!/bin/bash
python tools/step_trace.py --n 8192 > gpurun_out/r2x_trace_8192.txt 2>/dev/null
cat gpurun_out/r2x_trace_8192.txt | cut -c1-170 | head -70
