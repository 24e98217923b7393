# usage: tools/run_scale.sh <ngpus> <workload> <steps> <warmup> <tag>
N=$1; W=$2; S=$3; WU=$4; TAG=$5
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps $S --warmup $WU --workload $W > gpurun_out/scale_${TAG}.log 2>&1
tail -1 gpurun_out/scale_${TAG}.log | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print(d['n_gpus'], 'ms', round(d['ms_per_step'],1), 'e2e', d['e2e']['value'], {k: round(v,1) for k,v in d['stage_ms'].items()}, d['energy'])"
