#!/bin/bash
# round 3: the allocation-free batched entry (tests/test_zz_no_alloc.py) alone in a fresh process, then the whole GPU
# suite (which ends with it), then one bench line.  Logs go to gpurun_out/.
TAG=${1:-a}
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r3_noalloc_alone_$TAG.log 2>&1; echo "alone rc=$?"; tail -3 gpurun_out/r3_noalloc_alone_$TAG.log
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r3_gpu_suite_$TAG.log 2>&1; echo "suite rc=$?"; tail -3 gpurun_out/r3_gpu_suite_$TAG.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/r3_bench_$TAG.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/r3_bench_$TAG.log | cut -c1-400
