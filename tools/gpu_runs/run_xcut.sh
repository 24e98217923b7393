(timeout 300 python -m pytest tests -m gpu -x -q) > gpurun_out/pytest_chk.log 2>&1; tail -3 gpurun_out/pytest_chk.log
for wl in water3k mixed10k; do
  python tools/dump_grad.py $wl gpurun_out/g_${wl}_x60.npz
  for v in x55 x70; do DFTD4_B200_LIB=$PWD/dftd4_b200/lib/var/libdftd4_$v.so python tools/dump_grad.py $wl gpurun_out/g_${wl}_$v.npz; done
  python - <<PY
import numpy as np
a=np.load("gpurun_out/g_${wl}_x70.npz")
for v in ("x60","x55"):
    b=np.load("gpurun_out/g_${wl}_%s.npz"%v)
    print("${wl}",v,"dE rel %.2e"%(abs(a["energy"]-b["energy"])/abs(a["energy"])),"max dG %.2e"%np.max(np.abs(a["gradient"]-b["gradient"])),"max dS %.2e"%np.max(np.abs(a["virial"]-b["virial"])))
PY
done
B="timeout 300 python bench.py --steps 3 --warmup 3 --no-cpu-baseline"
brief() { python -c "
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
s=d['stage_ms']
print(sys.argv[1], 'ms', round(d['ms_per_step'],1), 'build', round(s['eeq_build'],1), 'solve', round(s['eeq_solve'],1), 'adj', round(s['eeq_adjoint'],1), 'E', d['energy'])
" $1; }
$B > gpurun_out/chk_x60.log 2>&1; brief gpurun_out/chk_x60.log
DFTD4_B200_LIB=$PWD/dftd4_b200/lib/var/libdftd4_x55.so $B > gpurun_out/chk_x55.log 2>&1; brief gpurun_out/chk_x55.log
