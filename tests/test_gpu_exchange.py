"""Result exchange over peer memory (include/rmpc.h rmpc_exchange_*, csrc/rmpc_exchange.cuh): two and three ranks played by
one process on one GPU — the regions are handed around as plain pointers instead of CUDA IPC handles, everything else
(stores into every rank's region, flags, the waiting warp, unpack, acknowledgements, the ring of `depth` slots and the wait
for a free slot) is the code the multi-process path runs.  The multi-process form is scripts/exchange_ranks.py (torchrun)."""
import os

import numpy as np
import pytest
import yaml

from conftest import SOLVERS
from robot_mpcs_b200 import scenarios as sc

pytestmark = pytest.mark.gpu

SDIR = "pointRobot_n3_005_H20_noSlack"


def _setup(world, per, depth, short_last=0):
    import torch
    from robot_mpcs_b200.parallel import PeerExchange
    from robot_mpcs_b200.planner.solver import Solver
    pm = yaml.safe_load(open(os.path.join(SOLVERS, SDIR, "paramMap.yaml")))
    solvers = [Solver.from_directory(os.path.join(SOLVERS, SDIR), device=0, max_batch=per) for _ in range(world)]
    regions = PeerExchange.alloc_regions(solvers[0], world, per, depth)
    xs = [PeerExchange(solvers[r], per, depth, rank=r, world=world, regions=regions) for r in range(world)]
    streams = [torch.cuda.Stream() for _ in range(world)]
    sizes = [per] * world
    sizes[-1] -= short_last
    return torch, pm, solvers, xs, streams, sizes, regions


def _teardown(solvers, xs, regions):
    for x in xs:
        x.close()
    for p in regions:
        solvers[0].lib.rmpc_ipc_free(0, p)


@pytest.mark.parametrize("world,short_last", [(2, 0), (3, 5)])
def test_every_rank_receives_every_ranks_results_over_more_steps_than_the_ring_has_slots(built, world, short_last):
    per, depth, steps = 96, 4, 11
    torch, pm, solvers, xs, streams, sizes, regions = _setup(world, per, depth, short_last)
    total = sum(sizes)
    pend, want = [], []
    for s in range(steps):
        data = sc.point_robot_sweep(pm, world * per, seed=100 + s)
        exp = []
        for r in range(world):   # every rank's solve + publish is enqueued before any collect of the step (one process plays all ranks)
            lo = r * per
            sl = slice(lo, lo + sizes[r])
            with torch.cuda.stream(streams[r]):
                out = solvers[r].solve_batch(torch.as_tensor(data["xinit"][sl], device="cuda:0"), torch.as_tensor(data["x0"][sl], device="cuda:0"),
                                             torch.as_tensor(data["params"][sl], device="cuda:0"), want_z=False, block=False)
                assert xs[r].publish(out["action"], out["exitflag"], out["iters"]) == s
            exp.append(out)
        want.append(exp)
        row = []
        for r in range(world):
            with torch.cuda.stream(streams[r]):
                row.append(xs[r].collect(s, total))
        pend.append(row)
    for s in range(steps):
        exp_a = torch.cat([o["action"] for o in want[s]]).cpu()
        exp_e = torch.cat([o["exitflag"] for o in want[s]]).cpu()
        exp_i = torch.cat([o["iters"] for o in want[s]]).cpu()
        for r in range(world):
            with torch.cuda.stream(streams[r]):
                a, e, i = pend[s][r].wait()
                a, e, i = a.cpu(), e.cpu(), i.cpu()
            if short_last:   # rows of the short shard beyond its size are not part of the batch
                keep = np.concatenate([np.arange(w * per, w * per + sizes[w]) for w in range(world)])
                full = pend[s][r]._out
                a, e, i = full[0].cpu()[keep], full[1].cpu()[keep], full[2].cpu()[keep]
            assert torch.equal(a, exp_a) and torch.equal(e, exp_e) and torch.equal(i, exp_i), (s, r)
        assert (exp_e == 1).all()
    torch.cuda.synchronize()
    assert [x.status() for x in xs] == [0] * world
    _teardown(solvers, xs, regions)


def test_exchange_refuses_misuse(built):
    torch, pm, solvers, xs, streams, sizes, regions = _setup(2, 32, 2)
    lib = solvers[0].lib
    a = torch.zeros(33, solvers[0].nu, dtype=torch.float64, device="cuda:0")
    z = torch.zeros(33, dtype=torch.int32, device="cuda:0")
    with pytest.raises(RuntimeError, match="rows_per_rank"):
        xs[0].publish(a, z, z)                      # more rows than the region holds
    with pytest.raises(RuntimeError, match="in order"):
        xs[0].collect(0, 64)                        # nothing published yet
    assert xs[0].publish(a[:32], z[:32], z[:32]) == 0 and xs[1].publish(a[:32], z[:32], z[:32]) == 0
    with pytest.raises(RuntimeError, match="in order"):
        xs[0].collect(1, 64)
    p0, p1 = xs[0].collect(0, 64), xs[1].collect(0, 64)
    assert float(p0.wait()[0].abs().sum()) == 0.0 and float(p1.wait()[0].abs().sum()) == 0.0
    assert lib.rmpc_exchange_region_bytes(0, 32, 2) == 0 and lib.rmpc_exchange_region_bytes(17, 32, 2) == 0
    assert not lib.rmpc_exchange_create(0, 2, 2, 32, 2, None)
    torch.cuda.synchronize()
    _teardown(solvers, xs, regions)
