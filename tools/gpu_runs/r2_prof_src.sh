#!/bin/bash
# usage: r2_prof_src.sh <tag> <kernel-regex> [env...]: source-level capture with the densest sampling (short kernels)
TAG=$1; KRE=$2; shift 2
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline"
env "$@" timeout 900 ncu --set full --warp-sampling-interval 0 --clock-control none --import-source on -k "regex:$KRE" -c ${NCAP:-1} -o gpurun_out/prof_$TAG $CMD > gpurun_out/ncu_full_$TAG.log 2>&1
tail -2 gpurun_out/ncu_full_$TAG.log
ncu -i gpurun_out/prof_$TAG.ncu-rep --page source --csv --print-source sass > gpurun_out/prof_${TAG}_sass.csv
ncu -i gpurun_out/prof_$TAG.ncu-rep --page source --csv --print-source cuda > gpurun_out/prof_${TAG}_cuda.csv 2>/dev/null
