#!/bin/bash
# tile ATM kernel: parity on the headline cuts, then A/B bench lines
mkdir -p gpurun_out
(timeout 900 python -m pytest tests/test_parity_gpu.py -m gpu -x -q -k "headline or seeded" ) > gpurun_out/tile_tests.log 2>&1; tail -5 gpurun_out/tile_tests.log
bash tools/gpu_runs/r2_ab.sh "$@"
