#!/usr/bin/env python
"""One FULL, unsampled evaluation of a workload by the CPU oracle (the reference's brute-force loops in C + OpenMP, EEQ
linear algebra through LAPACK like multicharge) on all host cores: the measured reference point bench.py's sampled CPU
baseline is checked against.  Prints one JSON line.  usage: cpu_full_reference.py <workload> [--check-gpu]

--check-gpu: the full-size cross-check.  The same structure goes through the public API on the GPU (DispersionModel.
get_dispersion) and the tool asserts the north_star tolerances against the CPU result: |dE| <= 1e-10 |E|, max |dg| and
max |dsigma| <= 1e-9 (Hartree/Bohr, Hartree), integer counts equal.  The differences are part of the JSON line."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dftd4_b200 import workloads  # noqa: E402
from oracle import pyoracle as O  # noqa: E402

args = [a for a in sys.argv[1:] if not a.startswith("--")]
name = args[0] if args else "water3k"
check_gpu = "--check-gpu" in sys.argv
O.build()
cores = O.set_num_threads(len(os.sched_getaffinity(0)))
lapack = O.use_lapack(True)
w = workloads.get(name)
p = O.load_rational_damping(w["method"], True)
t0 = time.perf_counter()
out = O.get_dispersion(w["numbers"], w["positions"], p, 0.0, w["lattice"], None, grad=True)
wall = time.perf_counter() - t0
check = None
if check_gpu:
    import numpy as np
    from dftd4_b200 import engine
    from dftd4_b200.interface import DampingParam, DispersionModel
    def self_images(cut):
        """Directed (i, i + t) entries within `cut`: atoms x non-zero lattice translations with |t| <= cut."""
        if w["lattice"] is None:
            return 0
        lat = np.asarray(w["lattice"], dtype=np.float64)
        r = range(-8, 9)
        t = np.array([[a, b, c] for a in r for b in r for c in r if (a, b, c) != (0, 0, 0)], dtype=np.float64) @ lat
        return int(len(w["numbers"]) * np.count_nonzero(np.einsum("ij,ij->i", t, t) <= cut * cut))

    model = DispersionModel(w["numbers"], w["positions"], lattice=w["lattice"])
    param = DampingParam(method=w["method"])
    gpu = model.get_dispersion(param, grad=True)
    ev = engine.ResidentEvaluation(model, param)
    ev.upload(); ev.compute(grad=True)
    cnt = ev.download(grad=True)["counts"]
    check = {"gpu_energy": float(gpu["energy"]), "rel_dE": abs(gpu["energy"] - out["energy"]) / abs(out["energy"]),
             "max_dg": float(np.max(np.abs(gpu["gradient"] - out["gradient"]))),
             "max_dsigma": float(np.max(np.abs(gpu["virial"] - out["sigma"]))),
             # the engine counts directed (i, j-image) entries, the oracle the reference's loop entries (i >= j, every
             # translation): directed = 2 x loop entries - self-image entries (an atom with its own images appears once
             # per translation in both); ordered triples t3 are counted alike
             "counts_equal": bool(int(cnt["t3"]) == int(out["counts"]["t3"]) and
                                  all(int(cnt[k]) == 2 * int(out["counts"][k]) - self_images(c) for k, c in
                                      (("pcn", 30.0), ("p2", 60.0), ("peeq", 25.0)))),
             "gpu_counts": cnt, "tolerances": {"rel_dE": 1e-10, "max_dg": 1e-9, "max_dsigma": 1e-9}}
    check["pass"] = bool(check["rel_dE"] <= 1e-10 and check["max_dg"] <= 1e-9 and check["max_dsigma"] <= 1e-9 and check["counts_equal"])
print(json.dumps({"gpu_check": check, "workload": w["name"], "natoms": int(len(w["numbers"])), "cores": cores, "lapack": lapack, "wall_s": wall,
                  "evals_per_s": 1.0 / wall, "energy": out["energy"], "counts": out["counts"], "stage_s": out["times"],
                  "eeq_s": O.eeq_times(), "kind": "port, full evaluation, unsampled"}))
if check is not None and not check["pass"]:
    sys.exit("full-size GPU/CPU cross-check FAILED: %s" % json.dumps(check))
