B="python bench.py --steps 3 --warmup 3 --no-cpu-baseline"
brief() { python -c "
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1], 'ms', round(d['ms_per_step'],1), 'atm', round(d['stage_ms']['atm'],1), 'frac', round(d['roofline']['frac'],3), 'E', d['energy'], 't3', d['counts']['t3'])
" $1; }
export D4CU_ATM_KERNEL=1
$B > gpurun_out/a1_default.log 2>&1; brief gpurun_out/a1_default.log
for v in "$@"; do DFTD4_B200_LIB=$PWD/dftd4_b200/lib/var/libdftd4_$v.so $B > gpurun_out/a1_$v.log 2>&1; brief gpurun_out/a1_$v.log; done
