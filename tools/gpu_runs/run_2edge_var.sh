#!/bin/bash
# two-edge ATM kernel variants: parity suite on the default build, then bench lines per variant library
mkdir -p gpurun_out
(timeout 600 python -m pytest tests -m gpu -x -q) > gpurun_out/2ev_tests.log 2>&1; tail -2 gpurun_out/2ev_tests.log
B="timeout 300 python bench.py --steps 4 --warmup 3 --no-cpu-baseline"
brief() { python -c "
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
s=d['stage_ms']
print(sys.argv[1], 'ms', round(d['ms_per_step'],1), 'atm', round(s['atm'],1), 'E', d['energy'])
" $1; }
$B > gpurun_out/2ev_default.log 2>&1; brief gpurun_out/2ev_default.log
for v in "$@"; do
  DFTD4_B200_LIB=$PWD/dftd4_b200/lib/var/libdftd4_$v.so $B > gpurun_out/2ev_$v.log 2>&1; brief gpurun_out/2ev_$v.log
done
