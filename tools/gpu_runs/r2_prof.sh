#!/bin/bash
# usage: r2_prof.sh <tag> <kernel-regex> [env assignments...]: one ncu --set full capture of the named kernel(s) on mixed10k
TAG=$1; KRE=$2; shift 2
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline"
env "$@" timeout 300 $CMD > gpurun_out/plain_$TAG.log 2>&1 || { tail -5 gpurun_out/plain_$TAG.log; exit 1; }
env "$@" timeout 900 ncu --set full --clock-control none --import-source on -k "regex:$KRE" -c ${NCAP:-1} -o gpurun_out/prof_$TAG $CMD > gpurun_out/ncu_full_$TAG.log 2>&1
tail -3 gpurun_out/ncu_full_$TAG.log
