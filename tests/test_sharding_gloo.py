"""Multi-GPU host logic on CPU: two gloo ranks take the row partition the engine uses (d4cu_partition: D4CU_ROW_BLOCK-atom blocks
dealt round-robin) and must cover every atom exactly once; the all-reduce of per-rank partial sums (what the engine
does with NCCL for energy / gradient / sigma) reproduces the single-rank result of a row-decomposable quantity."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, nat, out):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from dftd4_b200 import engine
    rows = engine.partition(nat, rank, world)
    owned = torch.zeros(nat, dtype=torch.int64)
    owned[torch.from_numpy(rows)] = 1
    dist.all_reduce(owned)
    # row-decomposable stand-in for the energy: sum over own rows of a pair function
    rng = np.random.default_rng(3)
    x = rng.uniform(0, 20, size=(nat, 3))
    d = np.linalg.norm(x[rows, None, :] - x[None, :, :], axis=-1)
    part = torch.tensor([np.sum(np.where(d > 0, 1.0 / (d ** 6 + 50.0), 0.0))], dtype=torch.float64)
    dist.all_reduce(part)
    if rank == 0:
        dfull = np.linalg.norm(x[:, None, :] - x[None, :, :], axis=-1)
        full = np.sum(np.where(dfull > 0, 1.0 / (dfull ** 6 + 50.0), 0.0))
        np.save(out, np.array([int(owned.min()), int(owned.max()), part.item(), full]))
    dist.destroy_process_group()


@pytest.mark.parametrize("nat", [1, 7, 8, 9, 63, 64, 65, 1000])
def test_partition_covers_rows_once_world2(tmp_path, d4lib, nat):
    out = str(tmp_path / "r.npy")
    mp.spawn(_worker, args=(2, _free_port(), nat, out), nprocs=2, join=True)
    lo, hi, part, full = np.load(out)
    assert lo == 1 and hi == 1
    assert abs(part - full) <= 1e-12 * abs(full)


def test_partition_rejects_bad_arguments(d4lib):
    from dftd4_b200 import engine
    with pytest.raises(ValueError):
        engine.partition(10, 2, 2)
    assert len(engine.partition(10, 0, 1)) == 10


@pytest.mark.parametrize("world", [1, 2, 3, 8])
def test_distributed_cholesky_column_ownership(d4lib, world):
    """cholesky_factor_dist: at every step each tile column of the trailing matrix is updated by exactly one rank, and
    that rank is the one that will factor the column when its turn comes (include/d4cu.h d4cu_chol_*)."""
    from dftd4_b200.library import lib
    nblk = 19
    for kb in range(nblk):
        rest = nblk - kb - 1
        seen = np.zeros(rest, dtype=int)
        for rank in range(world):
            t0 = lib.d4cu_chol_first_owned(kb, rank, world)
            assert 0 <= t0 < world
            for tj in range(t0, rest, world):
                seen[tj] += 1
                assert lib.d4cu_chol_column_owner(kb + 1 + tj, world) == rank
        assert np.all(seen == 1)


def test_workload_generators_are_pinned():
    """The synthetic BASELINE inputs are identified by the SHA-256 of their arrays: a changed generator, seed
    or numpy stream shows up here, not as a silently different benchmark."""
    import json
    import os
    from dftd4_b200 import workloads
    ref = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "workload_sha256.json")))
    for name in ("caffeine", "water1k", "water3k", "mixed10k"):
        w = workloads.get(name)
        assert len(w["numbers"]) == ref[name]["natoms"] and workloads.sha256(w) == ref[name]["sha256"], name
    w = workloads.conformers(2000)
    assert workloads.sha256(w) == ref["conformers-2000"]["sha256"]
