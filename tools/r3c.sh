#!/bin/bash
python tools/full_trace.py > gpurun_out/r3c_trace_full.txt 2>/dev/null
head -45 gpurun_out/r3c_trace_full.txt | cut -c1-160
