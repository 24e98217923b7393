#!/bin/bash
# usage: r2_multi.sh <ngpu>: multi-GPU parity tests, then one bench line at N GPUs
N=${1:-2}
mkdir -p gpurun_out
(timeout 1200 python -m pytest tests/test_multigpu.py -m gpu -x -q) > gpurun_out/multi_tests.log 2>&1; tail -4 gpurun_out/multi_tests.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29655 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/multi_bench_$N.log 2>&1
python - gpurun_out/multi_bench_$N.log <<'PY' || tail -20 gpurun_out/multi_bench_$N.log
import json, sys
d = [json.loads(l) for l in open(sys.argv[1]) if l.startswith('{')][-1]
print('N', d['n_gpus'], 'ms', round(d['ms_per_step'], 2), 'evals/s', round(d['value'], 3), 'e2e', round(d['e2e']['value'], 3),
      {k: round(v, 2) for k, v in d['stage_ms'].items()}, d.get('sharded_vs_single_gpu'), d.get('config_100k', {}).get('ms_per_step'))
PY
