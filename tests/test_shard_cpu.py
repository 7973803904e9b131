"""Multi-GPU host logic on CPU: slab arithmetic, and a world-size-2 gloo run in which each rank
processes its slab with a stand-in for the device call and the optional gather reassembles the
query order (the data path itself has no collective)."""
import os
import socket

import numpy as np
import pytest
import torch.multiprocessing as mp

from interpolate_unstructured_b200 import shard


@pytest.mark.parametrize("n,world", [(0, 1), (1, 2), (7, 2), (8, 2), (1_000_000_007, 8), (5, 8)])
def test_shard_range_partitions(n, world):
    edges = [shard.shard_range(n, k, world) for k in range(world)]
    assert edges[0][0] == 0 and edges[-1][1] == n
    for (a, b), (c, d) in zip(edges, edges[1:]):
        assert b == c and a <= b and c <= d
    sizes = [b - a for a, b in edges]
    assert max(sizes) - min(sizes) <= 1 and sizes == shard.shard_sizes(n, world)
    with pytest.raises(ValueError):
        shard.shard_range(n, world, world)


class _FakeGrid:
    """Stands in for api.Grid on CPU: a per-query function of the coordinates only, so sharded and
    unsharded evaluation must agree exactly."""

    def iu_interpolate_at(self, r, i_vars, i_cell=None):
        var = np.stack([r.sum(1) * v for v in i_vars], axis=1)
        return var, (np.floor(r[:, 0] * 1000).astype(np.int32) + 1)


def _worker(rank, world, port, n, out_dir):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    r = np.random.default_rng(7).random((n, 3))          # same global query array on every rank
    lo, hi, var, ic = shard.interpolate_sharded(_FakeGrid(), r, [1, 2, 3], rank, world)
    assert (lo, hi) == shard.shard_range(n, rank, world) and len(var) == hi - lo
    full_var = shard.gather_results(var, n)
    full_ic = shard.gather_results(ic, n)
    ref_var, ref_ic = _FakeGrid().iu_interpolate_at(r, [1, 2, 3])
    assert np.array_equal(full_var, ref_var) and np.array_equal(full_ic, ref_ic)
    open(os.path.join(out_dir, f"ok{rank}"), "w").close()
    dist.destroy_process_group()


def test_two_rank_gloo_shard_and_gather(tmp_path):
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    mp.spawn(_worker, args=(2, port, 10_001, str(tmp_path)), nprocs=2, join=True)
    assert sorted(os.listdir(tmp_path)) == ["ok0", "ok1"]
