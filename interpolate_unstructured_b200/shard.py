"""Query sharding for the multi-GPU path.

The hot path has no data dependence between queries and the mesh is read-only, so N GPUs run as N
independent replicas: the mesh is replicated (one ``Grid`` per process), the query range is cut into
N contiguous slabs, every rank writes its own output slab.  There is NO collective on the data
path.  ``gather_results`` is an optional epilogue (off the timed path) for callers that want every
rank to hold all results; it is the only function here that talks to ``torch.distributed``.
"""
from __future__ import annotations

import numpy as np


def shard_range(n, rank, world):
    """Half-open slab [lo, hi) of queries owned by `rank` out of `world`: contiguous, disjoint,
    covering 0..n, sizes differing by at most one (the first n % world ranks get the extra query)."""
    n, rank, world = int(n), int(rank), int(world)
    if world < 1 or not 0 <= rank < world or n < 0:
        raise ValueError("need 0 <= rank < world and n >= 0")
    q, rem = divmod(n, world)
    lo = rank * q + min(rank, rem)
    return lo, lo + q + (1 if rank < rem else 0)


def shard_sizes(n, world):
    return [hi - lo for lo, hi in (shard_range(n, k, world) for k in range(world))]


def interpolate_sharded(grid, r, i_vars, rank, world, i_cell=None):
    """This rank's share of a batched iu_interpolate_at over the global query array `r` (n, 3):
    returns (lo, hi, var_at_r[lo:hi], i_cell[lo:hi]).  `grid` is this rank's replica."""
    lo, hi = shard_range(len(r), rank, world)
    ic = None if i_cell is None else np.ascontiguousarray(i_cell[lo:hi], dtype=np.int32)
    var, ic = grid.iu_interpolate_at(np.ascontiguousarray(r[lo:hi]), i_vars, i_cell=ic)
    return lo, hi, var, ic


def gather_results(local, n_total, group=None):
    """All-gather of the per-rank result slabs (numpy, first axis = this rank's queries) into the
    full (n_total, ...) array in query order on every rank.  Works on any backend (gloo on CPU in
    the tests, nccl with CUDA tensors elsewhere); slabs may differ in length by one."""
    import torch
    import torch.distributed as dist
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    sizes = shard_sizes(n_total, world)
    if local.shape[0] != sizes[rank]:
        raise ValueError(f"rank {rank} holds {local.shape[0]} rows, expected {sizes[rank]}")
    pad = max(sizes)
    t = torch.zeros((pad,) + tuple(local.shape[1:]), dtype=torch.from_numpy(local[:0]).dtype)
    t[:sizes[rank]] = torch.from_numpy(np.ascontiguousarray(local))
    parts = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(parts, t, group=group)
    return np.concatenate([p[:s].numpy() for p, s in zip(parts, sizes)], axis=0)
