#!/bin/bash
# round 3, 8 GPUs: the bench line (with the 100k-atom point incl. its end-to-end time) and the batch50k line on 8 ranks.
TAG=${1:-a}; N=${2:-8}
mkdir -p gpurun_out
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $N --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r3_scale${N}_$TAG.log 2>&1; echo "bench rc=$?"
tail -1 gpurun_out/r3_scale${N}_$TAG.log | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print(d['n_gpus'], 'ms', round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],3), {k: round(v,1) for k,v in d['stage_ms'].items()}, d.get('sharded_vs_single_gpu'))
c=d.get('config_100k')
if c: print('100k ms', round(c['ms_per_step'],1), 'e2e ms', round(c['e2e_ms_per_step'],1), {k: round(v,1) for k,v in c['stage_ms'].items()}, c['energy'])" || tail -5 gpurun_out/r3_scale${N}_$TAG.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29532 bench.py --gpus $N --steps 5 --warmup 3 --workload batch50k > gpurun_out/r3_batch50k_n${N}_$TAG.log 2>&1; echo "batch rc=$?"; tail -1 gpurun_out/r3_batch50k_n${N}_$TAG.log | cut -c1-330
