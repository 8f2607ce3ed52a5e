#!/bin/bash
python tools/prof_hbm.py 3 > gpurun_out/r3j_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -k regex:"bias_act|image_logprob|noise_fill|clip_adam|grad_sumsq|returns_loss|kl_fwd|kl_bwd|pack_kernel" --csv --log-file gpurun_out/r3j_hbm.csv python tools/prof_hbm.py 3 > gpurun_out/r3j_ncu.log 2>&1
tail -n 2 gpurun_out/r3j_ncu.log; wc -l gpurun_out/r3j_hbm.csv
