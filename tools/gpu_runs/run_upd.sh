export DFTD4_B200_LIB=$PWD/dftd4_b200/lib/var/libdftd4_upd.so
(timeout 300 python -m pytest tests -m gpu -x -q) > gpurun_out/pytest_upd.log 2>&1; tail -3 gpurun_out/pytest_upd.log
B="timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline"
brief() { python -c "
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
s=d['stage_ms']
print(sys.argv[1], 'ms', round(d['ms_per_step'],1), 'solve', round(s['eeq_solve'],1), 'E', d['energy'])
" $1; }
$B > gpurun_out/upd_m10k.log 2>&1; brief gpurun_out/upd_m10k.log
$B --workload cluster20k > gpurun_out/upd_c20k.log 2>&1; brief gpurun_out/upd_c20k.log
unset DFTD4_B200_LIB
$B --workload cluster20k > gpurun_out/old_c20k.log 2>&1; brief gpurun_out/old_c20k.log
