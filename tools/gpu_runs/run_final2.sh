#!/bin/bash
# end-of-round: smoke, parity suite, official bench line, 100k atoms on one GPU, ncu launch list + full capture; usage run_final2.sh <tag>
TAG=$1
mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
(timeout 600 python -m pytest tests -m gpu -x -q) > gpurun_out/final_tests_$TAG.log 2>&1; tail -1 gpurun_out/final_tests_$TAG.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/final_mixed10k_$TAG.log 2>&1; tail -1 gpurun_out/final_mixed10k_$TAG.log | cut -c1-330
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/final_reference_$TAG.log 2>&1; tail -1 gpurun_out/final_reference_$TAG.log | cut -c1-300
timeout 600 python bench.py --workload mixed100k --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/final_mixed100k_$TAG.log 2>&1; tail -1 gpurun_out/final_mixed100k_$TAG.log | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('100k', round(d['ms_per_step'],1), {k: round(v,1) for k,v in d['stage_ms'].items()})"
NCAP=1 bash tools/gpu_profile.sh mixed10k $TAG "atm_two_edge_kernel"
