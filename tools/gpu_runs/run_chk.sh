(timeout 300 python -m pytest tests -m gpu -x -q) > gpurun_out/pytest_chk.log 2>&1; tail -4 gpurun_out/pytest_chk.log
B="timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline"
brief() { python -c "
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
s=d['stage_ms']
print(sys.argv[1], 'ms', round(d['ms_per_step'],1), 'atm', round(s['atm'],1), 'build', round(s['eeq_build'],1), 'solve', round(s['eeq_solve'],1), 'adj', round(s['eeq_adjoint'],1), 'E', d['energy'], 'launches', d['gpu_launches'])
" $1; }
$B > gpurun_out/chk_default.log 2>&1; brief gpurun_out/chk_default.log
D4CU_SUBST_STEPWISE=1 $B > gpurun_out/chk_step.log 2>&1; brief gpurun_out/chk_step.log
$B --workload cluster20k > gpurun_out/chk_c20k.log 2>&1; brief gpurun_out/chk_c20k.log
$B --workload water3k > gpurun_out/chk_w3k.log 2>&1; brief gpurun_out/chk_w3k.log
