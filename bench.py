#!/usr/bin/env python
"""bench.py -- D4 energy + gradient + virial evaluations per second (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload mixed10k|batch50k|...] [--impl ours|reference]

One process per GPU (torchrun sets RANK / LOCAL_RANK / WORLD_SIZE).  A "step" is one evaluation of the hot path
(src/dftd4/disp.f90:40-134 of the reference) on one synthetic structure of dftd4_b200.workloads.

  value          evaluations/s with the structure resident in HBM (d4cu_compute only; CUDA events on the engine stream)
  e2e            evaluations/s through the dftd4 C-API with HOST buffers: dftd4_update_structure + dftd4_get_dispersion,
                 i.e. host->device copy of coordinates and device->host copy of energy / gradient / virial per step
  roofline       dominant kernel (ATM triples, FP64 vector pipe): algorithmic flops / its CUDA-event time vs the measured
                 DFMA peak of this GPU (MEASURED_PEAKS.json has no FP64 entry, so the denominator is measured live)
  cpu_baseline   the CPU oracle (a port of the reference's loops) on the host cores, bounded sample, see _cpu_baseline

--impl reference times the reference algorithm on the host cores: the oracle port (the Fortran reference cannot be
built in this image: no Fortran compiler, mctc-lib / multicharge un-vendored).
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "D4 energy+gradient+virial evaluations/sec"
UNIT = "evals/s"
# algorithmic work per unit (DESIGN.md "work model"): flops per accepted interaction
FLOP_ATM_TRIPLE = 252.0   # atm.f90:347-425, derivatives
FLOP_PAIR_IMAGE = 88.0    # rational.f90:290-324, derivatives
# periodic EEQ (multicharge get_amat_3d / get_damat_3d real-space sums), per contributing image of a pair (alp r <= 6):
FLOP_EWALD_IMAGE = 13.0   # vector 3, r^2 5, sqrt 1, erfc 1, divide 1, weight + accumulate 2
FLOP_EWALD_DIMAGE = 38.0  # vector 3, r^2 5, sqrt 1, exp 1, erfc 1, radial factor 6, dG 6, dS 15


def _args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=os.environ.get("D4_BENCH_WORKLOAD", "mixed10k"))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


class ClockSampler:
    """SM clock / throttle reasons sampled with NVML while the timed region runs."""

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz, self._stop, self._t = [], set(), None, threading.Event(), None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:  # noqa: BLE001
            self.nv = None

    def _run(self):
        nv = self.nv
        names = {nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
                 nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap"}
        while not self._stop.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:  # noqa: BLE001
                pass
            self._stop.wait(0.05)

    def __enter__(self):
        if self.nv is not None:
            self._t = threading.Thread(target=self._run, daemon=True)
            self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self._t is not None:
            self._t.join()

    def summary(self):
        return {"sm_mhz": float(np.median(self.samples)) if self.samples else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons)}


def _host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def _sub_structures(name, w):
    """Two smaller structures of the same material for the EEQ stage of the CPU baseline (about 0.8k and 2.2k atoms)."""
    from dftd4_b200 import workloads
    if name.startswith("mixed"):
        return [workloads.mixed_cell(22.0), workloads.mixed_cell(30.0)]
    if name.startswith("cluster"):
        return [workloads.cluster(800), workloads.cluster(2200)]
    return []


def _cpu_baseline(name, w):
    """The reference algorithm (oracle port: the reference's brute-force loops in C + OpenMP, EEQ linear algebra through
    LAPACK dsytrf / dsytri / dgemm like multicharge) on the host cores, on a BOUNDED sample of workload `name`.

    The thread count is set explicitly to the cores this process may use (torchrun exports OMP_NUM_THREADS=1).
      pair + ATM   every s-th i-row of the two brute-force loops on the FULL structure; their cost is a sum over rows,
                   so the core-seconds spent in the sampled rows x s / cores is the time of a perfectly balanced run
      CN, C6, dCN  in full on the full structure
      EEQ          structures of up to 3 000 atoms: in full.  Larger: measured on two smaller structures of the same
                   material (N1 < N2 atoms); the O(N^2) parts (Ewald matrix + derivative matrices, dq/dr product with
                   dE/dq) are scaled by (N/N2)^2, the O(N^3) parts (dsytrf, dsytri, dgemm for dq/dr) by (N/N2)^3 --
                   the measured exponents between N1 and N2 are reported, and `eeq_extrapolated` says so.
    Returns (evals/s, description, cores, parts)."""
    from oracle import pyoracle as O
    cores = O.set_num_threads(_host_cores())
    lapack = O.use_lapack(True)
    n = len(w["numbers"])
    p = O.load_rational_damping(w["method"], True)
    if n <= 400:
        t0 = time.perf_counter()
        O.get_dispersion(w["numbers"], w["positions"], p, 0.0, w["lattice"], None, grad=True)
        total = time.perf_counter() - t0
        return 1.0 / total, "full %d-atom evaluation, unsampled" % n, cores, {"total_s": total, "eeq_extrapolated": False,
                                                                              "lapack": lapack}
    stride = max(1, n // max(24, 3 * cores))
    O.set_skip_eeq(True)
    try:
        out = O.get_dispersion(w["numbers"], w["positions"], p, 0.0, w["lattice"], None, grad=True, sample_stride=stride)
    finally:
        O.set_skip_eeq(False)
    t = out["times"]
    busy = O.row_busy()
    t_rows = (busy["disp2"] + busy["disp3"]) / cores * stride
    t_rest = t["cn"] + t["c6"] + t["c6_q0"] + t["dcn"]
    parts = {"pair_atm_s": t_rows, "cn_c6_s": t_rest, "rows_sampled": "1/%d" % stride, "lapack": lapack}

    def eeq_of(sub):
        ns = len(sub["numbers"])
        o = O.get_dispersion(sub["numbers"], sub["positions"], p, 0.0, sub["lattice"], None, grad=True, sample_stride=ns)
        e = O.eeq_times()
        return ns, e["build"] + e["dbuild"] + o["times"]["dq"], e["factor"] + e["inverse_gemm"]

    subs = _sub_structures(name, w) if n > 3000 else []
    if not subs:
        _, sq, cu = eeq_of(w)
        t_eeq = sq + cu
        parts.update(eeq_s=t_eeq, eeq_n2_s=sq, eeq_n3_s=cu, eeq_extrapolated=False)
        eeq_desc = "EEQ + dq/dr in full %.1fs" % t_eeq
    else:
        (n1, sq1, cu1), (n2, sq2, cu2) = eeq_of(subs[0]), eeq_of(subs[1])
        ex2 = float(np.log(sq2 / sq1) / np.log(n2 / n1))
        ex3 = float(np.log(cu2 / cu1) / np.log(n2 / n1))
        sq, cu = sq2 * (n / n2) ** 2, cu2 * (n / n2) ** 3
        t_eeq = sq + cu
        parts.update(eeq_s=t_eeq, eeq_n2_s=sq, eeq_n3_s=cu, eeq_extrapolated=True,
                     eeq_measured={"natoms": [n1, n2], "n2_parts_s": [sq1, sq2], "n3_parts_s": [cu1, cu2],
                                   "exponents": [ex2, ex3]})
        eeq_desc = ("EEQ + dq/dr measured at %d / %d atoms (matrix builds %.2f / %.2fs, exponent %.2f; dsytrf+dsytri+dgemm "
                    "%.2f / %.2fs, exponent %.2f), scaled (N/N2)^2 -> %.1fs and (N/N2)^3 -> %.1fs"
                    % (n1, n2, sq1, sq2, ex2, cu1, cu2, ex3, sq, cu))
    total = t_rows + t_rest + t_eeq
    parts["total_s"] = total
    desc = ("pair+ATM rows i%%%d==0 of the full %d-atom structure: core-seconds x%d / %d cores = %.1fs; CN+C6+dCN full %.1fs; %s"
            % (stride, n, stride, cores, t_rows, t_rest, eeq_desc))
    return 1.0 / total, desc, cores, parts


def _reference_arm(args, rank, world):
    """--impl reference: the reference algorithm on the host cores (oracle port), rank 0 only."""
    if rank != 0:
        return
    from dftd4_b200 import workloads
    from oracle import pyoracle as O
    O.build()
    w = workloads.get(args.workload)
    vals, desc, cores, parts, t_start, warm = [], "", 1, {}, time.perf_counter(), 0
    # every pass is a bounded sample (tens of seconds of CPU); the whole run is kept under ~3 minutes by dropping
    # warm-ups / steps that no longer fit, and "steps"/"warmup" report what actually ran
    for it in range(args.warmup + args.steps):
        if vals and time.perf_counter() - t_start > 150.0:
            break
        v, desc, cores, parts = _cpu_baseline(args.workload, w)
        if it < args.warmup and time.perf_counter() - t_start < 40.0:
            warm += 1
            continue
        vals.append(v)
    steps_requested, warmup_requested = args.steps, args.warmup
    args.warmup, args.steps = warm, len(vals)
    value = float(np.mean(vals))
    # "estimated": ms_per_step is the time of a FULL evaluation reconstructed from a bounded sample (sampled rows x stride,
    # EEQ extrapolated above 3 000 atoms), not the duration of a pass that ran; steps / warmup say what actually ran
    line = {"impl": "reference", "estimated": bool(len(w["numbers"]) > 400), "steps_requested": steps_requested,
            "warmup_requested": warmup_requested, "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 / value, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": w["name"], "natoms": int(len(w["numbers"])), "method": w["method"] + "-D4-ATM"},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": desc, "parts": parts},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


BATCH_METRIC = "batched D4 energy+gradient molecule evaluations/sec"
BATCH_UNIT = "molecules/s"
BATCH_NMOL = 50000


def _oracle_molecules(job):
    """Worker of the batch CPU baseline: one oracle call per molecule (what the reference API offers: one structure, one
    model, one get_dispersion per molecule), single-threaded; the pool supplies the parallelism over molecules."""
    from oracle import pyoracle as O
    numbers, positions, offsets, p = job
    O.set_num_threads(1)
    e = 0.0
    for m in range(len(offsets) - 1):
        lo, hi = offsets[m], offsets[m + 1]
        e += O.get_dispersion(numbers[lo:hi], positions[lo:hi], p, 0.0, grad=True)["energy"]
    return e


def _batch_cpu_baseline(w, nsample):
    """Molecules/s of the oracle port on the first `nsample` molecules, one process per host core."""
    import multiprocessing as mp
    from oracle import pyoracle as O
    O.build()
    cores = _host_cores()
    p = O.load_rational_damping(w["method"], True)
    off = w["offsets"]
    bounds = np.linspace(0, nsample, cores + 1).astype(int)
    jobs = []
    for a, b in zip(bounds[:-1], bounds[1:]):
        if b > a:
            lo, hi = off[a], off[b]
            jobs.append((w["numbers"][lo:hi], w["positions"][lo:hi], off[a:b + 1] - off[a], p))
    with mp.get_context("fork").Pool(len(jobs)) as pool:
        pool.map(_oracle_molecules, jobs[:1] * len(jobs))   # warm the workers (library load)
        t0 = time.perf_counter()
        pool.map(_oracle_molecules, jobs)
        dt = time.perf_counter() - t0
    return nsample / dt, cores, "first %d of %d molecules, one oracle call per molecule, %d worker processes" % (nsample, len(off) - 1, len(jobs))


def _batch_arm(args, rank, world, local):
    """--workload batch50k (BASELINE config 3): 50 000 drug-like conformers, B3LYP-D4-ATM energies + gradients + virials,
    ONE launch per batch.  A step is one pass over the whole batch; with N ranks every rank takes a contiguous 1/N of the
    molecules (independent objects: no collective on the data path)."""
    from dftd4_b200 import workloads
    w = workloads.conformers(BATCH_NMOL)
    nmol = BATCH_NMOL
    config = {"workload": "conformers-%d" % nmol, "molecules": nmol, "atoms": int(w["offsets"][-1]), "method": w["method"] + "-D4-ATM",
              "periodic": False, "cutoffs_bohr": [60.0, 40.0, 30.0], "parallelism": "molecules x%d" % world}
    if args.impl == "reference":
        if rank != 0:
            return
        nsample = 20000
        vals = []
        for _ in range(max(1, min(args.steps, 3))):
            v, cores, desc = _batch_cpu_baseline(w, nsample)
            vals.append(v)
        value = float(np.mean(vals))
        print(json.dumps({"impl": "reference", "metric": BATCH_METRIC, "value": value, "unit": BATCH_UNIT, "n_gpus": args.gpus,
                          "steps": len(vals), "steps_requested": args.steps, "warmup": 1, "estimated": True,
                          "ms_per_step": 1e3 * nmol / value, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                          "dtype": "f64", "data": "synthetic", "config": config,
                          "cpu_baseline": {"value": value, "unit": BATCH_UNIT, "cores": cores, "kind": "port", "sample": desc},
                          "e2e": {"value": value, "unit": BATCH_UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                          "gpu_launches": 0}), flush=True)
        return

    import torch
    import torch.distributed as dist

    from dftd4_b200 import engine
    from dftd4_b200.batch import get_dispersion_batch
    from dftd4_b200.interface import DampingParam
    if engine.device_count() < 1:
        raise SystemExit("bench.py: no CUDA device -- the D4 engine has no CPU fallback")
    torch.cuda.set_device(local)
    os.environ.setdefault("D4CU_DEVICE", str(local))
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    # this rank's molecules; inputs in pinned host memory (numpy views over pinned torch tensors)
    m0, m1 = rank * nmol // world, (rank + 1) * nmol // world
    a0, a1 = int(w["offsets"][m0]), int(w["offsets"][m1])
    pos_pin = torch.empty((a1 - a0, 3), dtype=torch.float64).pin_memory()
    num_pin = torch.empty((a1 - a0,), dtype=torch.int32).pin_memory()
    off_pin = torch.empty((m1 - m0 + 1,), dtype=torch.int64).pin_memory()
    pos, num, off = pos_pin.numpy(), num_pin.numpy(), off_pin.numpy()
    pos[:] = w["positions"][a0:a1]; num[:] = w["numbers"][a0:a1]; off[:] = w["offsets"][m0:m1 + 1] - a0
    p = DampingParam(method=w["method"]).as_tuple()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        out = get_dispersion_batch(off, num, pos, p, grad=True)
    assert np.all(out["status"] == 0)
    kern_ms, wall = 0.0, 0.0
    barrier()
    with ClockSampler(local) as clk:
        for _ in range(args.steps):
            flush.zero_()
            barrier()
            t0 = time.perf_counter()
            out = get_dispersion_batch(off, num, pos, p, grad=True)   # H2D, one kernel launch, D2H of E / gradient / virial
            wall += time.perf_counter() - t0
            kern_ms += out["kernel_ms"]                                 # CUDA events around the launch, engine stream
    t = torch.tensor([kern_ms, wall], dtype=torch.float64, device="cuda")
    esum = torch.tensor([float(out["energy"].sum())], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(esum)
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    ms_per_step = float(t[0].item()) / args.steps
    value = nmol / (ms_per_step * 1e-3)
    e2e = nmol * args.steps / float(t[1].item())
    # algorithmic work of the batch kernel: every pair and triple of a molecule is inside the cutoffs (molecules span
    # ~15 bohr), plus the in-shared-memory Cholesky of the (n+1)^2 EEQ system
    sizes = np.diff(w["offsets"]).astype(np.float64)
    flops = float(np.sum(sizes * (sizes - 1) / 2 * FLOP_PAIR_IMAGE + sizes * (sizes - 1) * (sizes - 2) / 6 * FLOP_ATM_TRIPLE
                         + sizes ** 3 / 3.0))
    peaks = engine.measure_fp64_peak(local)
    clocks = clk.summary()
    roof = {"bound": "fp64", "kernel": "batch_kernel", "achieved": flops / (ms_per_step * 1e-3) / 1e12, "peak": peaks["dfma_tflops"],
            "unit": "TFLOP/s", "n_gpus": world,
            # dram__bytes_read.sum + dram__bytes_write.sum of one batch_kernel launch over the 50 000 molecules, from the
            # committed ncu --set full capture (72.7 MB + 26.2 MB; the inputs are 70 MB, the outputs 64 MB of which the L2
            # still held most at the end of the launch): a constant from that profile, not measured in this run
            "traffic": 98.93e6 if world == 1 else None,
            "traffic_source": "profiles/r03_batch50k_ncu_full_a.txt" if world == 1 else None,
            "flops_per_batch": flops, "work_model": "pairs x %.0f + triples x %.0f + n^3/3 (EEQ Cholesky) flop" % (FLOP_PAIR_IMAGE, FLOP_ATM_TRIPLE),
            "peak_source": "measured live: register-resident DFMA chain on all SMs (d4cu_measure_fp64_peak)"}
    roof["frac"] = roof["achieved"] / (roof["peak"] * world) if roof["peak"] else None
    if clocks.get("sm_mhz"):
        roof["peak_nominal_clock"] = 148 * 64 * 2 * clocks["sm_mhz"] * 1e6 / 1e12
    h2d = int(pos.nbytes + num.nbytes + off.nbytes)
    d2h = int(out["energy"].nbytes + out["gradient"].nbytes + out["virial"].nbytes + out["status"].nbytes)
    line = {"metric": BATCH_METRIC, "value": value, "unit": BATCH_UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic", "config": dict(config, l2="flushed between timed steps (256 MiB memset, untimed)"),
            "e2e": {"value": e2e, "unit": BATCH_UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "note": "per rank; d4cu_batch_numbers with host buffers: species model build, H2D, one launch, D2H"},
            "gpu_launches": args.steps, "energy_sum": float(esum.item()), "clocks": clocks, "roofline": roof}
    if not args.no_cpu_baseline and world == 1:
        v, cores, desc = _batch_cpu_baseline(w, 20000)
        line["cpu_baseline"] = {"value": v, "unit": BATCH_UNIT, "cores": cores, "kind": "port", "sample": desc}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    args = _args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.workload == "batch50k":
        _batch_arm(args, rank, world, local)
        return
    if args.impl == "reference":
        _reference_arm(args, rank, world)
        return

    import torch
    import torch.distributed as dist

    from dftd4_b200 import engine, workloads
    from dftd4_b200.interface import DampingParam, DispersionModel

    if engine.device_count() < 1:
        raise SystemExit("bench.py: no CUDA device -- the D4 engine has no CPU fallback")
    torch.cuda.set_device(local)
    os.environ.setdefault("D4CU_DEVICE", str(local))
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    w = workloads.get(args.workload)
    n = len(w["numbers"])
    # pinned host staging for the e2e leg (numpy views over pinned torch tensors)
    pos_pin = torch.empty((n, 3), dtype=torch.float64).pin_memory()
    pos = pos_pin.numpy()
    pos[:] = w["positions"]
    model = DispersionModel(w["numbers"], pos, lattice=w["lattice"])
    param = DampingParam(method=w["method"])
    ev = engine.ResidentEvaluation(model, param)
    if world > 1:
        uid = [engine.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        ev.set_comm(rank, world, uid[0])

    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- resident leg: inputs in HBM, d4cu_compute only -----------------------------------------------------------
    ev.upload()
    for _ in range(args.warmup):
        ev.compute(grad=True)
    stage_sum, dev_ms, launches = {}, 0.0, 0
    barrier()
    with ClockSampler(local) as clk:
        t_wall = time.perf_counter()
        for _ in range(args.steps):
            flush.zero_()          # L2 flush between timed iterations (untimed: the step time is the engine's events)
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()     # the ranks start a step together: a late starter would wait inside the first collective
            ev.compute(grad=True)
            st, nl = ev.timing()
            dev_ms += st["total"]
            launches += nl
            for k, v in st.items():
                stage_sum[k] = stage_sum.get(k, 0.0) + v
        barrier()
        t_wall = time.perf_counter() - t_wall
    res = ev.download(grad=True)
    tmax = torch.tensor([dev_ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    ms_per_step = float(tmax.item()) / args.steps
    value = 1e3 / ms_per_step
    # per-rank time of the dominant (three-body) stage: with N > 1 the ranks carry different shares of it (rank 0 also
    # holds the EEQ system), the roofline line counts GPU-seconds over all ranks
    atm_all = torch.zeros(world, dtype=torch.float64, device="cuda")
    atm_all[rank] = stage_sum.get("atm", 0.0) / args.steps
    if world > 1:
        dist.all_reduce(atm_all)
    atm_ms_ranks = [float(x) for x in atm_all.tolist()]

    # ---- e2e leg: host buffers through the dftd4 C-API ------------------------------------------------------------
    for _ in range(min(args.warmup, 2)):
        model.update(pos)
        model.get_dispersion(param, grad=True)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        model.update(pos)                       # dftd4_update_structure: new coordinates from (pinned) host memory
        out = model.get_dispersion(param, grad=True)   # H2D, all kernels, D2H of E / gradient / sigma
    barrier()
    t_e2e = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t_e2e, op=dist.ReduceOp.MAX)
    e2e = args.steps / float(t_e2e.item())
    assert abs(out["energy"] - res["energy"]) <= 1e-12 * abs(res["energy"])

    # ---- N > 1: the sharded gradient / virial against one evaluation on rank 0 alone (untimed) -------------------------
    shard_check = None
    if world > 1:
        if rank == 0:
            solo = DispersionModel(w["numbers"], pos, lattice=w["lattice"]).get_dispersion(param, grad=True)
            shard_check = {"max_abs_dgradient": float(np.max(np.abs(solo["gradient"] - res["gradient"]))),
                           "max_abs_dvirial": float(np.max(np.abs(solo["virial"] - res["virial"]))),
                           "rel_denergy": abs(solo["energy"] - res["energy"]) / abs(solo["energy"])}
            assert shard_check["max_abs_dgradient"] <= 1e-11 and shard_check["max_abs_dvirial"] <= 1e-11, shard_check
        barrier()

    # ---- --gpus 8 only: the 100k-atom point of BASELINE's metric, 2 steps, reported under an extra key ---------------------
    extra_100k = None
    if world == 8 and args.workload == "mixed10k" and not os.environ.get("D4_BENCH_NO_100K"):
        w2 = workloads.get("mixed100k")
        m2 = DispersionModel(w2["numbers"], w2["positions"], lattice=w2["lattice"])
        ev2 = engine.ResidentEvaluation(m2, param)
        uid = [engine.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        ev2.set_comm(rank, world, uid[0])
        ev2.upload()
        ev2.compute(grad=True)                      # warm-up (allocations)
        barrier()
        t100, st100 = 0.0, {}
        for _ in range(2):
            ev2.compute(grad=True)
            st100, _nl = ev2.timing()
            t100 += st100["total"]
        tm = torch.tensor([t100 / 2], dtype=torch.float64, device="cuda")
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
        r100 = ev2.download(grad=True)
        # end to end at 100k atoms (structure ingest): host coordinates through dftd4_update_structure
        # (wrap + cold-fusion sweep on the host, H2D) + dftd4_get_dispersion (all kernels, D2H), one timed evaluation
        barrier()
        t0 = time.perf_counter()
        m2.update(w2["positions"])
        o100 = m2.get_dispersion(param, grad=True)
        barrier()
        te = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
        assert abs(o100["energy"] - r100["energy"]) <= 1e-12 * abs(r100["energy"])
        e2e_100k_ms = 1e3 * float(te.item())
        extra_100k = {"workload": w2["name"], "natoms": int(len(w2["numbers"])), "steps": 2, "ms_per_step": float(tm.item()),
                      "evals_per_s": 1e3 / float(tm.item()), "e2e_ms_per_step": e2e_100k_ms,
                      "e2e_minus_resident_ms": e2e_100k_ms - float(tm.item()), "stage_ms": st100, "energy": r100["energy"],
                      "counts": r100["counts"]}
        del ev2, m2
        barrier()

    if rank == 0:
        stages = {k: v / args.steps for k, v in stage_sum.items()}
        peaks = engine.measure_fp64_peak(local)
        c = res["counts"]
        tri = c["t3"] / 6.0                      # unordered accepted triangles (reference loop entries x triple_scale)
        # all ranks' triangles / the slowest rank's stage time (the stage's critical path); the mean over the ranks
        # (GPU-seconds / N) is reported beside it as `achieved_mean_rank`
        atm_s = max(atm_ms_ranks) * 1e-3
        atm_mean_s = sum(atm_ms_ranks) / world * 1e-3
        pair_s = stages["pair"] * 1e-3
        dom = max(("atm", "pair", "eeq_solve", "eeq_adjoint", "lists"), key=lambda k: stages[k])
        roof = {"bound": "fp64", "kernel": "atm_stream_kernel", "achieved": tri * FLOP_ATM_TRIPLE / atm_s / 1e12 if atm_s > 0 else 0.0,
                "achieved_mean_rank": tri * FLOP_ATM_TRIPLE / atm_mean_s / 1e12 if atm_mean_s > 0 else 0.0,
                "peak": peaks["dfma_tflops"], "unit": "TFLOP/s",
                # 148 SMs x 64 FP64 lanes x 2 flop x SM clock: the arithmetic peak at the recorded clock, stated next
                # to the chain-measured one (which is what `frac` divides by)
                "peak_nominal_clock": None,
                # dram__bytes_read.sum + dram__bytes_write.sum of one atm_stream_kernel launch, taken from the committed
                # ncu --set full capture of this workload (profiles/r02_mixed10k_ncu_full_b.txt: 2.738 GB + 0.322 GB: the
                # 64-byte records and C6 rows streamed by the bulk copies, the per-entry accumulators written); a constant
                # from that profile, not measured in this run, and not captured for the other workloads
                "traffic": 3.0594e9 if (w["name"].startswith("mixed-cell-9819") and world == 1) else None,
                "traffic_source": "profiles/r02_mixed10k_ncu_full_b.txt" if (w["name"].startswith("mixed-cell-9819") and world == 1) else None,
                "peak_source": "measured live: register-resident DFMA chain on all SMs (d4cu_measure_fp64_peak); "
                               "MEASURED_PEAKS.json has no FP64 entry",
                "dmma_peak_tflops": peaks["dmma_tflops"],
                "pair_kernel_tflops": (c["p2"] / 2.0) * FLOP_PAIR_IMAGE / pair_s / 1e12 if pair_s > 0 else 0.0,
                "largest_stage": dom}
        # the other rooflines north_star names: DMMA rate of the Cholesky (ld^3/3 flop over the factor+substitution stage,
        # i.e. a lower bound for the trailing-update kernel) and HBM rate of the CN pass (32-B image points walked)
        ld = ((n + 127) // 128) * 128
        if stages["eeq_solve"] > 0:
            roof["eeq_factor_tflops"] = ld ** 3 / 3.0 / (stages["eeq_solve"] * 1e-3) / 1e12
        if w["lattice"] is not None and stages["eeq_build"] > 0:
            # periodic EEQ: algorithmic flops of the real-space sums over the exactly counted contributing images / stage
            # time (the stages also hold the rank-128 reciprocal update resp. the substitution + reciprocal pass)
            img = ev.ewald_images()
            roof["ewald_images"] = int(img)
            roof["eeq_build_tflops"] = img * FLOP_EWALD_IMAGE / (stages["eeq_build"] * 1e-3) / 1e12
            roof["eeq_adjoint_tflops"] = img * FLOP_EWALD_DIMAGE / (stages["eeq_adjoint"] * 1e-3) / 1e12
        if stages["cn"] > 0:   # (0.2 ms at 10k atoms: latency-bound at this size, quoted for completeness)
            roof["cn_pass_gbs"] = c["pcn"] * 32.0 / (stages["cn"] * 1e-3) / 1e9
        try:
            with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
                roof["hbm_peak_gbs"] = json.load(f).get("hbm_gbs")   # driver-measured copy bandwidth (denominator of cn_pass_gbs)
        except Exception:  # noqa: BLE001
            pass
        # `achieved` is the whole job's algorithmic rate (all ranks' triangles / max-over-ranks time); the fraction is per GPU
        roof["frac"] = roof["achieved"] / (roof["peak"] * world) if roof["peak"] else None
        roof["n_gpus"] = world
        if world > 1:
            roof["atm_ms_per_rank"] = atm_ms_ranks
        clocks = clk.summary()
        if clocks.get("sm_mhz"):
            roof["peak_nominal_clock"] = 148 * 64 * 2 * clocks["sm_mhz"] * 1e6 / 1e12
            roof["frac_nominal_clock"] = roof["achieved"] / (roof["peak_nominal_clock"] * world)
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "f64", "data": "synthetic",
                "config": {"workload": w["name"], "natoms": int(n), "method": w["method"] + "-D4-ATM",
                           "periodic": w["lattice"] is not None, "cutoffs_bohr": [60.0, 40.0, 30.0],
                           "l2": "flushed between timed steps (256 MiB memset, untimed)",
                           "parallelism": "rows x%d" % world},
                "e2e": {"value": e2e, "unit": UNIT, "h2d_bytes_per_step": int(n * 24 + n * 4 + (72 if w["lattice"] is not None else 0)),
                        "d2h_bytes_per_step": int(8 + n * 24 + 72 + 80)},
                "gpu_launches": int(launches), "stage_ms": stages, "counts": c, "energy": res["energy"],
                "wall_ms_per_step_incl_flush": 1e3 * t_wall / args.steps, "clocks": clocks, "roofline": roof}
        if shard_check is not None:
            line["sharded_vs_single_gpu"] = shard_check
        if extra_100k is not None:
            line["config_100k"] = extra_100k
        if not args.no_cpu_baseline and world == 1:
            from oracle import pyoracle as O
            O.build()
            v, desc, cores, parts = _cpu_baseline(args.workload, w)
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": desc, "parts": parts}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
