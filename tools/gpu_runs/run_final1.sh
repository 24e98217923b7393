#!/bin/bash
# final single-GPU lines of the round: usage run_final1.sh <tag>
TAG=$1
mkdir -p gpurun_out
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/final_mixed10k_$TAG.log 2>&1; tail -1 gpurun_out/final_mixed10k_$TAG.log | cut -c1-400
for w in water3k cluster20k caffeine; do
  timeout 600 python bench.py --steps 5 --warmup 3 --workload $w --no-cpu-baseline > gpurun_out/final_${w}_$TAG.log 2>&1
  tail -1 gpurun_out/final_${w}_$TAG.log | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$w', round(d['ms_per_step'],2), 'ms', d['value'], d['unit'], 'e2e', d['e2e']['value'])" || tail -3 gpurun_out/final_${w}_$TAG.log
done
