#!/bin/bash
# round 3: final 1-GPU bench record, ncu launch list of the same command, one full capture of pair_kernel.
TAG=${1:-a}
mkdir -p gpurun_out
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/r3_bench_final_$TAG.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/r3_bench_final_$TAG.log | cut -c1-200
CMD="python bench.py --steps 2 --warmup 1 --no-cpu-baseline"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/r3_launches_$TAG.csv $CMD > gpurun_out/r3_ncu_list_$TAG.log 2>&1; echo "list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:pair_kernel -c 1 -o gpurun_out/r3_prof_pair_$TAG $CMD > gpurun_out/r3_ncu_pair_$TAG.log 2>&1; echo "ncu rc=$?"
ncu -i gpurun_out/r3_prof_pair_$TAG.ncu-rep --page raw --csv > gpurun_out/r3_prof_pair_${TAG}_raw.csv 2>&1
ncu -i gpurun_out/r3_prof_pair_$TAG.ncu-rep --page details > gpurun_out/r3_prof_pair_${TAG}_details.txt 2>&1
