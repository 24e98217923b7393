#!/usr/bin/env python
"""BASELINE config 3: batched drug-like conformers (~50 atoms each), B3LYP-D4 energies + gradients, one launch per batch.

    python tools/bench_batch.py [--nmol 50000] [--steps 5] [--warmup 3]

Prints one JSON line: molecule evaluations per second for (a) the batch kernel alone (CUDA events), (b) the whole call
with host buffers (H2D + kernel + D2H), (c) a loop of per-molecule calls through the dftd4 API mirror on a sample, and
(d) the CPU oracle (reference algorithm) on a sample using all host cores."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--nmol", type=int, default=50000)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--no-cpu", action="store_true")
    a = ap.parse_args()
    from dftd4_b200 import workloads
    from dftd4_b200.batch import get_dispersion_batch
    from dftd4_b200.interface import DampingParam, DispersionModel
    w = workloads.conformers(a.nmol)
    param = DampingParam(method=w["method"])
    p = param.as_tuple()
    for _ in range(a.warmup):
        out = get_dispersion_batch(w["offsets"], w["numbers"], w["positions"], p, grad=True)
    kern, wall = [], []
    for _ in range(a.steps):
        t0 = time.perf_counter()
        out = get_dispersion_batch(w["offsets"], w["numbers"], w["positions"], p, grad=True)
        wall.append(time.perf_counter() - t0)
        kern.append(out["kernel_ms"] * 1e-3)
    assert np.all(out["status"] == 0)
    nloop = min(200, a.nmol)
    t0 = time.perf_counter()
    for m in range(nloop):
        lo, hi = w["offsets"][m], w["offsets"][m + 1]
        one = DispersionModel(w["numbers"][lo:hi], w["positions"][lo:hi]).get_dispersion(param, grad=True)
    t_loop = (time.perf_counter() - t0) / nloop
    assert abs(one["energy"] - out["energy"][nloop - 1]) <= 1e-12 * abs(one["energy"])
    line = {"metric": "batched D4 energy+gradient molecule evaluations/sec", "unit": "molecules/s", "nmol": a.nmol,
            "atoms": int(w["offsets"][-1]), "method": w["method"] + "-D4-ATM", "steps": a.steps, "warmup": a.warmup,
            "kernel": a.nmol / float(np.mean(kern)), "kernel_ms": 1e3 * float(np.mean(kern)),
            "e2e": a.nmol / float(np.mean(wall)), "e2e_ms": 1e3 * float(np.mean(wall)),
            "per_molecule_api_loop": 1.0 / t_loop, "gpu_launches_per_batch": 1,
            "h2d_bytes": int(w["positions"].nbytes + w["numbers"].nbytes + w["offsets"].nbytes),
            "d2h_bytes": int(out["energy"].nbytes + out["gradient"].nbytes + out["virial"].nbytes)}
    if not a.no_cpu:
        from oracle import pyoracle as O
        O.build()
        po = O.load_rational_damping(w["method"], True)
        ns = min(300, a.nmol)
        t0 = time.perf_counter()
        for m in range(ns):
            lo, hi = w["offsets"][m], w["offsets"][m + 1]
            O.get_dispersion(w["numbers"][lo:hi], w["positions"][lo:hi], po, 0.0, grad=True)
        line["cpu_baseline"] = {"value": ns / (time.perf_counter() - t0), "unit": "molecules/s", "cores": O.num_threads(),
                                "kind": "port", "sample": "first %d molecules, one oracle call each" % ns}
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
