#!/bin/bash
# two-edge ATM kernel: GPU parity suite, then A/B against the one-edge kernel (D4CU_ATM_SCHEME=1)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/2e_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/2e_tests.log
tail -5 gpurun_out/2e_tests.log
timeout 300 python bench.py --steps 5 --warmup 3 > gpurun_out/2e_bench_new.log 2>&1; tail -1 gpurun_out/2e_bench_new.log | cut -c1-1500
D4CU_ATM_SCHEME=1 timeout 300 python bench.py --steps 5 --warmup 3 > gpurun_out/2e_bench_old.log 2>&1; tail -1 gpurun_out/2e_bench_old.log | cut -c1-600
