#!/bin/bash
# Run on the GPU box (via gpurun): plain bench, then the ncu launch list and one full capture of the top kernels.
# usage: tools/gpu_profile.sh <workload> <tag> [kernel-regex]
W=${1:-water3k}; TAG=${2:-r1}; KRE=${3:-atm_kernel|eeq_grad_kernel|eeq_build_3d_kernel|pair_kernel}
CMD="python bench.py --workload $W --steps 1 --warmup 1 --no-cpu-baseline"
mkdir -p gpurun_out
$CMD > gpurun_out/plain_$TAG.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/launches_$TAG.csv $CMD > gpurun_out/ncu_launch_$TAG.log 2>&1
$CMD > gpurun_out/plain2_$TAG.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k "regex:$KRE" -c ${NCAP:-4} -o gpurun_out/prof_$TAG $CMD > gpurun_out/ncu_full_$TAG.log 2>&1
tail -3 gpurun_out/ncu_full_$TAG.log
