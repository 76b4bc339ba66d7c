"""Multi-GPU sharding of the interpolation path: one process per
GPU (``torch.distributed``), query points split into contiguous shards exactly
like the reference's ``parallel_for`` splits them across threads
(cxx/include/pyinterp/parallel_for.hpp:46-65), the grid replicated once with a
single broadcast (NCCL over NVLink on GPUs, gloo in the CPU tests).  There is
no data-path collective: results stay sharded; the only cross-rank reduction is
the bounds-error "first offending point" (a min over ranks).
"""
from __future__ import annotations

import numpy as np

from . import _abi


def init(n_gpus: int = 0) -> int:
    """Single-process multi-GPU: from now on every host-array call of
    ``bivariate`` / ``trivariate`` / ``quadrivariate`` is split into contiguous
    shards over the first ``n_gpus`` CUDA devices (0: all of them) INSIDE the
    one C-ABI call (``pyi_dist_init``): the grid handle is replicated per
    device over NVLink peer copies, each shard runs its own pipelined copies
    and kernels, the result array and the bounds-error report are those of a
    one-device call.  This is the drop-in shape of the reference, whose
    ``parallel_for`` splits one call's points over the workers of one process
    (cxx/include/pyinterp/parallel_for.hpp:30-78).  Returns the device count in
    effect."""
    n = int(_abi.load().pyi_dist_init(int(n_gpus)))
    if n <= 0:
        raise _abi.PyiError("dist.init: no CUDA device (pyinterp_b200 has no CPU fallback)")
    return n


def size() -> int:
    """Device count host-array calls are sharded over (1 until ``init``)."""
    return int(_abi.load().pyi_dist_size())


def shard_bounds(n: int, world_size: int, rank: int) -> tuple[int, int]:
    """[start, end) of `rank`: n // W points each, the first n % W ranks take
    one more (parallel_for.hpp:46-53)."""
    if world_size <= 0 or not (0 <= rank < world_size):
        raise ValueError("invalid rank / world size")
    chunk, rem = divmod(int(n), world_size)
    start = rank * chunk + min(rank, rem)
    return start, start + chunk + (1 if rank < rem else 0)


def broadcast_grid_array(array, src: int = 0, group=None):
    """Replicate the grid values from rank `src` (one broadcast per grid handle).
    `array` is a torch tensor (CUDA for NCCL, CPU for gloo) with the final
    shape / dtype on every rank; returns it filled."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.broadcast(array, src=src, group=group)
    return array


def first_out_of_bounds(local, group=None):
    """Global bounds-error report: the lowest offending point index over all
    ranks (and its axis), as the reference reports the lowest chunk's first
    failure (parallel_for.hpp:71-77).  `local` is None or (global_index, axis)."""
    import torch
    import torch.distributed as dist
    big = np.iinfo(np.int64).max
    key = big if local is None else (int(local[0]) << 2) | int(local[1])
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        backend = dist.get_backend(group)
        dev = torch.device("cuda", torch.cuda.current_device()) if backend == "nccl" else torch.device("cpu")
        t = torch.tensor([key], dtype=torch.int64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MIN, group=group)
        key = int(t.item())
    if key == big:
        return None
    return key >> 2, key & 3
