"""Scenario sharding across ranks on CPU: world_size 2, gloo.  The solve itself needs a GPU; here
each rank fabricates its shard's (action, exitflag, iters) from the global scenario index, and the gathered result
must be the whole batch in order on every rank."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from robot_mpcs_b200.parallel import gather_results, shard_bounds


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, B, nu, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lo, hi = shard_bounds(B, rank, world)
    idx = torch.arange(lo, hi, dtype=torch.float64)
    action = idx[:, None] * 10 + torch.arange(nu, dtype=torch.float64)[None, :]
    flag = torch.where(idx.to(torch.int64) % 3 == 0, -7, 1).to(torch.int32)
    iters = (idx.to(torch.int64) % 17).to(torch.int32)
    a, f, it = gather_results(action, flag, iters, B)
    # the same in flight: two gathers issued before either is waited for (a throughput loop waits once, at the end)
    p1 = gather_results(action, flag, iters, B, async_op=True)
    p2 = gather_results(action + 1.0, flag, iters, B, async_op=True)
    a2, _, _ = p2.wait()
    a1, f1, it1 = p1.wait()
    same = torch.equal(a1, a) and torch.equal(f1, f) and torch.equal(it1, it) and torch.equal(a2, a + 1.0)
    ref = torch.arange(B, dtype=torch.float64)
    ok = (a.shape == (B, nu) and torch.equal(a, ref[:, None] * 10 + torch.arange(nu, dtype=torch.float64)[None, :])
          and torch.equal(f, torch.where(ref.to(torch.int64) % 3 == 0, -7, 1).to(torch.int32))
          and torch.equal(it, (ref.to(torch.int64) % 17).to(torch.int32)))
    out[rank] = int(ok and same)
    dist.destroy_process_group()


@pytest.mark.parametrize("B", [10, 7, 1])
def test_gather_world2(B):
    world, nu = 2, 3
    out = mp.Manager().dict()
    mp.spawn(_worker, args=(world, _free_port(), B, nu, out), nprocs=world, join=True)
    assert dict(out) == {0: 1, 1: 1}


def test_shard_bounds_cover():
    for B in (1, 5, 8, 4096, 4097):
        for world in (1, 2, 3, 8):
            spans = [shard_bounds(B, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == B
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            assert max(hi - lo for lo, hi in spans) == (B + world - 1) // world
