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
