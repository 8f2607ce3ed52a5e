#!/bin/bash
python tools/step_trace.py > gpurun_out/r3o_trace_2500.txt 2>/dev/null
head -48 gpurun_out/r3o_trace_2500.txt | cut -c1-150
