"""Multi-GPU parity (needs >= 2 CUDA devices; skipped otherwise): gradient and virial of the row-sharded evaluation
(world = 2, NCCL) equal the single-GPU evaluation -- in EEQ owner mode (rank 0 alone factors while rank 1 works on
three-body rows) and with the distributed (column-cyclic) Cholesky factorisation -- and a rank that owns no rows still
reaches every collective."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

_SCRIPT = r"""
import json, os, sys
import numpy as np
sys.path.insert(0, sys.argv[1])
import torch, torch.distributed as dist
from dftd4_b200 import engine, workloads
from dftd4_b200.interface import DampingParam, DispersionModel
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
os.environ["D4CU_DEVICE"] = str(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
case = sys.argv[3]
if case == "tiny":      # 6 atoms on 2 ranks with 8-row blocks: rank 1 owns NO rows and must still reach every collective
    w = dict(numbers=np.array([8, 1, 1] * 2, dtype=np.int32), lattice=None, method="pbe",
             positions=np.array([[3.0 * (k // 3), 1.5 * (k % 3), 0.7 * (k % 2)] for k in range(6)]))
else:                   # periodic mixed cell with Si, ~1.2k atoms
    w = workloads.mixed_cell(24.0)
def run(comm):
    model = DispersionModel(w["numbers"], w["positions"], lattice=w["lattice"])
    ev = engine.ResidentEvaluation(model, DampingParam(method=w["method"]))
    if comm:
        uid = [engine.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        ev.set_comm(rank, world, uid[0])
    ev.upload(); ev.compute(grad=True)
    return ev.download(grad=True)
if case == "props":
    # owner mode (rank 0 alone solves the EEQ system): properties and the disp3 pair list on EVERY rank.  The charges,
    # polarisabilities and C6 a rank returns must be the solved ones, not whatever its own q buffer held.
    os.environ["D4CU_DEVICE"] = str(local)
    model = DispersionModel(w["numbers"], w["positions"], lattice=w["lattice"])
    ev = engine.ResidentEvaluation(model, DampingParam(method=w["method"]))
    uid = [engine.nccl_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(uid, src=0)
    ev.set_comm(rank, world, uid[0])
    props = model.get_properties()                       # collective: both ranks call it
    ev.upload(); ev.compute(grad=True)
    i3, j3, t3 = ev.pair_list(2)
    parts = [None] * world
    dist.all_gather_object(parts, np.stack([i3, j3, t3], axis=1))
    ref_model = DispersionModel(w["numbers"], w["positions"], lattice=w["lattice"])
    ref = ref_model.get_properties()                     # no communicator: one GPU
    ev1 = engine.ResidentEvaluation(ref_model, DampingParam(method=w["method"]))
    ev1.upload(); ev1.compute(grad=True)
    one = np.stack(ev1.pair_list(2), axis=1)
    both = np.concatenate(parts, axis=0)
    key = lambda a: a[np.lexsort((a[:, 2], a[:, 1], a[:, 0]))]
    diffs = {k: float(np.max(np.abs(np.asarray(props[k]) - np.asarray(ref[k])))) for k in
             ("partial charges", "coordination numbers", "polarizabilities", "c6 coefficients") if k in props}
    assert len(diffs) >= 3, list(props)
    json.dump({"diffs": diffs, "pairs_equal": bool(both.shape == one.shape and np.array_equal(key(both), key(one))),
               "npairs": int(one.shape[0])}, open(sys.argv[2] + ".%d" % rank, "w"))
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0)
sharded = run(True)
dist.barrier()
if rank == 0:
    single = run(False)
    json.dump({"e": [sharded["energy"], single["energy"]],
               "dg": float(np.max(np.abs(sharded["gradient"] - single["gradient"]))),
               "ds": float(np.max(np.abs(sharded["virial"] - single["virial"]))),
               "counts": [sharded["counts"], single["counts"]]}, open(sys.argv[2], "w"))
dist.barrier()
dist.destroy_process_group()
"""


def _ndev():
    from dftd4_b200 import engine
    return engine.device_count()


@pytest.mark.parametrize("case,chol_dist", [("mixed", "0"), ("mixed", "1"), ("tiny", "0")])
def test_sharded_evaluation_equals_single_gpu(d4lib, tmp_path, case, chol_dist):
    if _ndev() < 2:
        pytest.skip("needs two CUDA devices")
    out = tmp_path / "res.json"
    env = dict(os.environ, D4CU_CHOL_DIST=chol_dist)
    script = tmp_path / "shard.py"
    script.write_text(_SCRIPT)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29617", str(script), ROOT, str(out), case]
    res = subprocess.run(cmd, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=900)
    assert res.returncode == 0, res.stdout[-3000:]
    r = json.load(open(out))
    assert abs(r["e"][0] - r["e"][1]) <= 1e-13 * abs(r["e"][1])
    assert r["dg"] <= 1e-12 and r["ds"] <= 1e-12
    assert r["counts"][0] == r["counts"][1]


def test_owner_mode_properties_and_pair_list_on_every_rank(d4lib, tmp_path):
    """EEQ owner mode keeps the solved charges on rank 0; d4cu_properties must hand them to the other ranks before it
    derives weights, polarisabilities and C6 from them, and d4cu_pair_list(which = 2) must map the three-body rows
    through the row set that mode dealt (union over the ranks = the single-GPU list)."""
    if _ndev() < 2:
        pytest.skip("needs two CUDA devices")
    out = tmp_path / "props.json"
    script = tmp_path / "props.py"
    script.write_text(_SCRIPT)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29618", str(script), ROOT, str(out), "props"]
    res = subprocess.run(cmd, env=dict(os.environ, D4CU_CHOL_DIST="0"), stdout=subprocess.PIPE, stderr=subprocess.STDOUT,
                         text=True, timeout=900)
    assert res.returncode == 0, res.stdout[-3000:]
    for rank in range(2):
        r = json.load(open(str(out) + ".%d" % rank))
        assert all(v <= 1e-12 for v in r["diffs"].values()), (rank, r["diffs"])
        assert r["pairs_equal"] and r["npairs"] > 0, rank
