#!/bin/bash
# usage: r2_ab.sh [-t] [-w workload] "<env assignments>" ...: one short bench line per environment setting (A/B runs);
# -t runs the GPU parity tests first.  Output: gpurun_out/ab_<n>.log + one summary line each.
mkdir -p gpurun_out
WL=mixed10k
while getopts "tw:" o; do
  case $o in
    t) (timeout 1500 python -m pytest tests -m gpu -x -q) > gpurun_out/ab_tests.log 2>&1; tail -3 gpurun_out/ab_tests.log;;
    w) WL=$OPTARG;;
  esac
done
shift $((OPTIND - 1))
n=0
for cfg in "$@"; do
  n=$((n + 1))
  env $cfg timeout 600 python bench.py --steps 4 --warmup 3 --no-cpu-baseline --workload $WL > gpurun_out/ab_$n.log 2>&1
  python - "$cfg" gpurun_out/ab_$n.log <<'PY' || tail -5 gpurun_out/ab_$n.log
import json, sys
d = json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
s = d['stage_ms']
print('[%s]' % sys.argv[1], 'ms', round(d['ms_per_step'], 1), ' '.join('%s %.1f' % (k, v) for k, v in s.items() if v >= 0.05),
      'E', d['energy'], 't3', d['counts']['t3'], 'frac', round(d['roofline']['frac'], 4))
PY
done
