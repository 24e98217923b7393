#!/bin/bash
# usage: r2_final.sh <tag> [full3k]: the round-2 measurement set on one B200 -- bench lines, reference arm, ncu launch list
# and one full capture of the dominant kernel (each after its plain run exited 0); `full3k` adds the unsampled CPU
# evaluation of the 3 000-atom water box (minutes of host time).
TAG=${1:-a}
mkdir -p gpurun_out
nproc > gpurun_out/host_$TAG.txt; lscpu | grep -E "Model name|^CPU\(s\)|Thread|Socket" >> gpurun_out/host_$TAG.txt; free -g | head -2 >> gpurun_out/host_$TAG.txt
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_mixed10k_$TAG.log 2>&1; tail -1 gpurun_out/bench_mixed10k_$TAG.log | cut -c1-200
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference_$TAG.log 2>&1; tail -1 gpurun_out/bench_reference_$TAG.log | cut -c1-200
for wl in water3k cluster20k caffeine; do
  timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --workload $wl > gpurun_out/bench_${wl}_$TAG.log 2>&1; tail -1 gpurun_out/bench_${wl}_$TAG.log | cut -c1-160
done
CMD="python bench.py --steps 2 --warmup 1 --no-cpu-baseline"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/launches_$TAG.csv $CMD > gpurun_out/ncu_list_$TAG.log 2>&1; tail -1 gpurun_out/ncu_list_$TAG.log | cut -c1-120
timeout 900 ncu --set full --clock-control none --import-source on -k regex:atm_stream -c 1 -o gpurun_out/prof_atm_$TAG $CMD > gpurun_out/ncu_full_$TAG.log 2>&1
ncu -i gpurun_out/prof_atm_$TAG.ncu-rep --page raw --csv > gpurun_out/prof_atm_${TAG}_raw.csv
ncu -i gpurun_out/prof_atm_$TAG.ncu-rep --page source --csv --print-source sass > gpurun_out/prof_atm_${TAG}_sass.csv
ncu -i gpurun_out/prof_atm_$TAG.ncu-rep --page details > gpurun_out/prof_atm_${TAG}_details.txt
if [ "$2" = "full3k" ]; then
  timeout 3000 python tools/cpu_full_reference.py water3k > gpurun_out/cpu_full_water3k_$TAG.json 2> gpurun_out/cpu_full_water3k_$TAG.err; tail -1 gpurun_out/cpu_full_water3k_$TAG.json | cut -c1-300
fi
