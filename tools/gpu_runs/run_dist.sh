N=$1
run() { # tag workload envs...
  TAG=$1; WL=$2; shift 2
  env "$@" timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519 bench.py --gpus $N --steps 3 --warmup 3 --workload $WL --no-cpu-baseline > gpurun_out/dist_$TAG.log 2>&1
  tail -1 gpurun_out/dist_$TAG.log | python -c "
import json,sys
d=json.loads(sys.stdin.read())
s=d['stage_ms']
print('$TAG', d['n_gpus'], 'ms', round(d['ms_per_step'],1), 'solve', round(s['eeq_solve'],1), 'build', round(s['eeq_build'],1), 'atm', round(s['atm'],1), 'adj', round(s['eeq_adjoint'],1), 'E', d['energy'])" || tail -5 gpurun_out/dist_$TAG.log
}
run m10k_dist mixed10k D4CU_CHOL_DIST=1
run m10k_single mixed10k D4CU_CHOL_DIST=0
run c20k_dist cluster20k D4CU_CHOL_DIST=1
run c20k_single cluster20k D4CU_CHOL_DIST=0
