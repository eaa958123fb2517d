"""N>1 host logic on CPU: world_size-2 (and 3, ragged) `gloo` runs of the shard/solve/gather driver with a
stand-in solve function, compared with the single-process result.  The CUDA solve itself is covered by the
`-m gpu` tests; here only partitioning and result collection are under test."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from g1_locomotion_b200 import sharding


def _fake_inputs(lo, hi, N=4, nu=6):
    idx = torch.arange(lo, hi, dtype=torch.float64)
    x0 = idx[:, None] * torch.ones(1, 13, dtype=torch.float64)
    load = (idx[:, None, None] + 1.0) * torch.ones(1, N, nu, dtype=torch.float64)
    return x0, load


def _fake_solve(x0, load):
    """Deterministic per-problem function standing in for BatchedMPC.update_batch."""
    u = load * 2.0 + x0[:, None, :1]
    x = torch.cumsum(x0[:, None, :].expand(-1, load.shape[1], -1), dim=1)
    return u, x, dict(status=torch.zeros(x0.shape[0], dtype=torch.int32))


def _worker(rank, world, port, total, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        got, u, x, info = sharding.solve_sharded(_fake_solve, _fake_inputs, total)
        lo, hi = sharding.shard_range(total, rank, world)
        assert u.shape[0] == hi - lo
        torch.save(got, os.path.join(out_dir, f"rank{rank}.pt"))
    finally:
        dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def test_shard_ranges_partition_the_batch():
    for total in (0, 1, 7, 1024, 1048576):
        for world in (1, 2, 3, 4, 8):
            rs = [sharding.shard_range(total, r, world) for r in range(world)]
            assert rs[0][0] == 0 and rs[-1][1] == total
            assert all(rs[i][1] == rs[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in rs]
            assert max(sizes) - min(sizes) <= 1 and sizes == sharding.shard_sizes(total, world)
    with pytest.raises(ValueError):
        sharding.shard_range(8, 2, 2)


@pytest.mark.parametrize("world,total", [(2, 64), (3, 10)])
def test_gloo_sharded_solve_equals_single_process(tmp_path, world, total):
    want, _, _, _ = sharding.solve_sharded(_fake_solve, _fake_inputs, total)     # no process group: whole batch
    assert want.shape == (total, 6 + 13)
    mp.spawn(_worker, args=(world, _free_port(), total, str(tmp_path)), nprocs=world, join=True)
    for r in range(world):
        got = torch.load(os.path.join(str(tmp_path), f"rank{r}.pt"))
        assert torch.equal(got, want), r
