#!/bin/bash
# final 8-GPU lines: usage run_scale8f.sh <tag>
T=$1
run() { # tag workload steps warmup
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 --steps $3 --warmup $4 --workload $2 --no-cpu-baseline > gpurun_out/scale8_$1.log 2>&1
  tail -1 gpurun_out/scale8_$1.log | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('$1', d['n_gpus'], 'ms', round(d['ms_per_step'],1), 'e2e', round(d['e2e']['value'],3), {k: round(v,1) for k,v in d['stage_ms'].items()}, 'E', d['energy'])" || tail -5 gpurun_out/scale8_$1.log
}
run mixed10k_$T mixed10k 5 3
run cluster20k_$T cluster20k 3 3
run mixed100k_$T mixed100k 1 1
