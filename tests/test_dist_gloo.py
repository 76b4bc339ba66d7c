"""Multi-rank host logic of the path on CPU (gloo, world_size 2): contiguous
point shards like the reference's parallel_for
(cxx/include/pyinterp/parallel_for.hpp:46-65), one broadcast of the grid per
handle, and the bounds-error reduction (lowest offending point over ranks).
There is no data-path collective to test: results stay sharded."""
import importlib
import os
import socket

import numpy as np
import pytest


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, n, tmp):
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, root)
    import torch
    import torch.distributed as dist
    pd = importlib.import_module("pangeo-pyinterp_b200.dist")
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        # 1. the grid is replicated once from rank 0
        ref = torch.arange(12 * 7, dtype=torch.float64).reshape(12, 7)
        arr = ref.clone() if rank == 0 else torch.zeros_like(ref)
        pd.broadcast_grid_array(arr, src=0)
        assert torch.equal(arr, ref)
        # 2. contiguous shards
        b, e = pd.shard_bounds(n, world, rank)
        local = np.arange(b, e, dtype=np.int64)
        np.save(os.path.join(tmp, f"shard{rank}.npy"), local)
        # 3. bounds error: rank 1 sees an earlier point of its shard on axis 1,
        # rank 0 a later point of its own shard on axis 0 -> rank 0's global
        # index is lower, so it wins
        mine = (b + 5, 0) if rank == 0 else (b + 1, 1)
        got = pd.first_out_of_bounds(mine)
        assert got == (5, 0), got
        # only one rank fails
        got = pd.first_out_of_bounds((b + 2, 3) if rank == 1 else None)
        assert got == (pd.shard_bounds(n, world, 1)[0] + 2, 3), got
        # nobody fails
        assert pd.first_out_of_bounds(None) is None
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(180)
def test_two_ranks_gloo(tmp_path):
    import torch.multiprocessing as mp
    n, world = 1001, 2
    mp.spawn(_worker, args=(world, _free_port(), n, str(tmp_path)), nprocs=world, join=True)
    shards = [np.load(tmp_path / f"shard{r}.npy") for r in range(world)]
    assert np.array_equal(np.concatenate(shards), np.arange(n))
    assert abs(len(shards[0]) - len(shards[1])) <= 1 and len(shards[0]) >= len(shards[1])


@pytest.mark.parametrize("n,world", [(0, 1), (1, 8), (7, 8), (8, 8), (1000, 3), (10**10, 8)])
def test_shard_bounds_matches_parallel_for(n, world):
    pd = importlib.import_module("pangeo-pyinterp_b200.dist")
    chunk, rem = divmod(n, world)
    pos = 0
    for r in range(world):
        b, e = pd.shard_bounds(n, world, r)
        assert b == pos and e - b == chunk + (1 if r < rem else 0)
        pos = e
    assert pos == n
    with pytest.raises(ValueError):
        pd.shard_bounds(n, world, world)
