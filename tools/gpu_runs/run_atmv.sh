B="timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline"
brief() { python -c "
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[1], 'ms', round(d['ms_per_step'],1), 'atm', round(d['stage_ms']['atm'],1), 'E', d['energy'], 't3', d['counts']['t3'])
" $1; }
for v in "$@"; do DFTD4_B200_LIB=$PWD/dftd4_b200/lib/var/libdftd4_$v.so $B > gpurun_out/atmv_$v.log 2>&1; brief gpurun_out/atmv_$v.log; done
