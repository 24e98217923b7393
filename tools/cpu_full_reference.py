#!/usr/bin/env python
"""One FULL, unsampled evaluation of a workload by the CPU oracle (the reference's brute-force loops in C + OpenMP, EEQ
linear algebra through LAPACK like multicharge) on all host cores: the measured reference point bench.py's sampled CPU
baseline is checked against.  Prints one JSON line.  usage: cpu_full_reference.py <workload>"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dftd4_b200 import workloads  # noqa: E402
from oracle import pyoracle as O  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "water3k"
O.build()
cores = O.set_num_threads(len(os.sched_getaffinity(0)))
lapack = O.use_lapack(True)
w = workloads.get(name)
p = O.load_rational_damping(w["method"], True)
t0 = time.perf_counter()
out = O.get_dispersion(w["numbers"], w["positions"], p, 0.0, w["lattice"], None, grad=True)
wall = time.perf_counter() - t0
print(json.dumps({"workload": w["name"], "natoms": int(len(w["numbers"])), "cores": cores, "lapack": lapack, "wall_s": wall,
                  "evals_per_s": 1.0 / wall, "energy": out["energy"], "counts": out["counts"], "stage_s": out["times"],
                  "eeq_s": O.eeq_times(), "kind": "port, full evaluation, unsampled"}))
