#!/bin/bash
# round 3: BASELINE config 3 through bench.py (--workload batch50k): our arm, the reference arm, an ncu launch list and one
# full capture of batch_kernel (each after its plain run exited 0); then the full-size water3k GPU/CPU cross-check.
TAG=${1:-a}
mkdir -p gpurun_out
timeout 900 python bench.py --workload batch50k --steps 5 --warmup 3 > gpurun_out/r3_bench_batch50k_$TAG.log 2>&1; echo "batch rc=$?"; tail -1 gpurun_out/r3_bench_batch50k_$TAG.log | cut -c1-600
timeout 600 python bench.py --workload batch50k --impl reference --steps 2 --warmup 1 > gpurun_out/r3_bench_batch50k_ref_$TAG.log 2>&1; echo "ref rc=$?"; tail -1 gpurun_out/r3_bench_batch50k_ref_$TAG.log | cut -c1-300
CMD="python bench.py --workload batch50k --steps 1 --warmup 1 --no-cpu-baseline"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:batch_kernel -c 1 -o gpurun_out/r3_prof_batch_$TAG $CMD > gpurun_out/r3_ncu_batch_$TAG.log 2>&1; echo "ncu rc=$?"
ncu -i gpurun_out/r3_prof_batch_$TAG.ncu-rep --page details > gpurun_out/r3_prof_batch_${TAG}_details.txt 2>&1
ncu -i gpurun_out/r3_prof_batch_$TAG.ncu-rep --page raw --csv > gpurun_out/r3_prof_batch_${TAG}_raw.csv 2>&1
if [ "$2" = "full3k" ]; then
  timeout 1200 python tools/cpu_full_reference.py water3k --check-gpu > gpurun_out/r3_fullcheck_water3k.json 2> gpurun_out/r3_fullcheck_water3k.err; echo "fullcheck rc=$?"; cut -c1-420 gpurun_out/r3_fullcheck_water3k.json
fi
