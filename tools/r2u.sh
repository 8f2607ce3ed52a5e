#!/bin/bash
python -m pytest tests/test_noise_gpu.py -m gpu -q -x 2>&1 | tail -25 > gpurun_out/r2u_noise.log; cat gpurun_out/r2u_noise.log | tail -25 | cut -c1-300
python -m pytest tests -m gpu -q 2>&1 | tail -40 > gpurun_out/r2u_tests.log
grep -E "passed|failed|FAILED|^E  " gpurun_out/r2u_tests.log | head -30 | cut -c1-300
bash tools/r2t.sh
