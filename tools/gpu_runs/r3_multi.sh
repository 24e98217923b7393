#!/bin/bash
# round 3, N GPUs (default 2): multi-GPU parity tests (log kept for profiles/), then the bench line on N GPUs.
N=${1:-2}; TAG=${2:-a}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_multigpu.py -m gpu -x -q -rs > gpurun_out/r3_multigpu_tests_$TAG.log 2>&1; echo "multigpu tests rc=$?"; tail -5 gpurun_out/r3_multigpu_tests_$TAG.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29655 bench.py --gpus $N --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r3_bench_n${N}_$TAG.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/r3_bench_n${N}_$TAG.log | cut -c1-300
