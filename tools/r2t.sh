#!/bin/bash
# round-2 final bench lines (one GPU) + ncu launch list + ncu --set full of the four rollout kernels
set -o pipefail
for spec in "walker_large bf16" "config4 bf16" "pong bf16" "scaled bf16" "walker fp32"; do
  set -- $spec
  python bench.py --steps 20 --warmup 5 --workload $1 --precision $2 --no-cpu-baseline 2> gpurun_out/r2t_$1_$2.err | grep "^{" > gpurun_out/r2t_bench_$1_$2.json
  python - "$1" "$2" <<'PY'
import json,sys
try:
    d=json.loads(open(f'gpurun_out/r2t_bench_{sys.argv[1]}_{sys.argv[2]}.json').read())
    print(sys.argv[1], sys.argv[2], round(d['ms_per_step'],3), round(d['value']/1e6,2), 'e2e', round(d['e2e']['value']/1e6,2),
          'ref_cuda', (d.get('reference_cuda') or {}).get('speedup_vs_best_case'), 't_full', (d.get('t_full') or {}).get('ms'),
          'full', (d.get('full_update') or {}).get('ms'), 'roof', d['roofline'].get('frac'))
except Exception as e: print(sys.argv[1], 'FAILED', e)
PY
done
CMD="python bench.py --steps 2 --warmup 1 --no-graph --no-cpu-baseline --no-reference-cuda --no-full --no-sustained --no-secondary"
$CMD > gpurun_out/r2t_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2t_bf16_launches.csv $CMD > gpurun_out/r2t_ncu1.log 2>&1
python tools/prof_r2.py 2 > gpurun_out/r2t_plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:rollout_.*kernel -s 4 -c 4 -o gpurun_out/r2t_rollouts python tools/prof_r2.py 2 > gpurun_out/r2t_ncu2.log 2>&1
tail -n 2 gpurun_out/r2t_ncu1.log; tail -n 2 gpurun_out/r2t_ncu2.log; ls -la gpurun_out/*.ncu-rep
