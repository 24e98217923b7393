run() { # tag workload steps warmup [env...]
  TAG=$1; WL=$2; S=$3; WU=$4; shift 4
  env "$@" timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 --steps $S --warmup $WU --workload $WL --no-cpu-baseline > gpurun_out/scale8_$TAG.log 2>&1
  tail -1 gpurun_out/scale8_$TAG.log | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('$TAG', d['n_gpus'], 'ms', round(d['ms_per_step'],1), 'e2e', round(d['e2e']['value'],3), {k: round(v,1) for k,v in d['stage_ms'].items()}, 'E', d['energy'])" || tail -5 gpurun_out/scale8_$TAG.log
}
run mixed10k_dist mixed10k 5 3 D4CU_CHOL_DIST=1
run mixed10k mixed10k 5 3 D4CU_CHOL_DIST=0
run cluster20k cluster20k 3 3 D4CU_X=0
run mixed100k mixed100k 1 1 D4CU_X=0
