#!/bin/bash
# round 2: streaming ATM kernels vs the round-1 kernel; parity suite first, then bench lines per configuration
mkdir -p gpurun_out
(timeout 900 python -m pytest tests -m gpu -x -q) > gpurun_out/r2a_tests.log 2>&1; tail -3 gpurun_out/r2a_tests.log
B="timeout 300 python bench.py --steps 4 --warmup 3 --no-cpu-baseline"
brief() { python -c "
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
s=d['stage_ms']
print(sys.argv[1], 'ms', round(d['ms_per_step'],1), 'atm', round(s['atm'],1), 'lists', round(s['lists'],1), 'E', d['energy'], 't3', d['counts']['t3'])
" $1 || tail -5 $1; }
D4CU_ATM_SCHEME=1 $B > gpurun_out/r2a_v1.log 2>&1; brief gpurun_out/r2a_v1.log
$B > gpurun_out/r2a_s0.log 2>&1; brief gpurun_out/r2a_s0.log
D4CU_SORT_FINE=2 $B > gpurun_out/r2a_s2.log 2>&1; brief gpurun_out/r2a_s2.log
D4CU_SORT_FINE=3 $B > gpurun_out/r2a_s3.log 2>&1; brief gpurun_out/r2a_s3.log
D4CU_SORT_FINE=2 D4CU_ATM_KC=8 $B > gpurun_out/r2a_s2k8.log 2>&1; brief gpurun_out/r2a_s2k8.log
D4CU_SORT_FINE=2 $B --workload water3k > gpurun_out/r2a_w3k.log 2>&1; brief gpurun_out/r2a_w3k.log
