#!/usr/bin/env python
"""Multi-process check of the peer-memory result exchange (include/rmpc.h rmpc_exchange_*): under torchrun, one rank per GPU,
every rank solves its shard on rotating streams, publishes, collects, and compares every step with an NCCL all_gather of the
same tensors.  Prints one JSON line on rank 0.
usage: python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 scripts/exchange_ranks.py"""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch, torch.distributed as dist, yaml
from robot_mpcs_b200 import scenarios as sc
from robot_mpcs_b200.parallel import PeerExchange, gather_results, init_process_group
from robot_mpcs_b200.planner.solver import Solver

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
init_process_group("nccl", device_id=dev)
sdir = os.path.join(ROOT, "robot_mpcs_b200", "solvers", "boxer_n3_01_H30_noSlack")
pm = yaml.safe_load(open(os.path.join(sdir, "paramMap.yaml")))
from robot_mpcs_b200.parallel import shard_bounds
steps, depth = int(os.environ.get("XR_STEPS", "30")), int(os.environ.get("XR_DEPTH", "4"))
total = 1024 * world - 7                              # the last shard is short
per = -(-total // world)
lo, hi = shard_bounds(total, rank, world)
B = hi - lo
solver = Solver.from_directory(sdir, device=local, max_batch=per)
x = PeerExchange(solver, per, depth)
streams = [torch.cuda.Stream(dev) for _ in range(3)]
sets = []
for i in range(3):
    s = sc.boxer_radial(pm, per, seed=50 * rank + i)
    sets.append({k: torch.as_tensor(v[:B], device=dev) for k, v in s.items() if k in ("xinit", "x0", "params")})
pend, outs = [], []
torch.cuda.synchronize(); dist.barrier()
t = time.perf_counter()
for i in range(steps):
    with torch.cuda.stream(streams[i % 3]):
        d = sets[i % 3]
        r = solver.solve_batch(d["xinit"], d["x0"], d["params"], want_z=False)
        pend.append(x.collect(x.publish(r["action"], r["exitflag"], r["iters"]), total))
        outs.append(r)
got = [p.wait() for p in pend]
torch.cuda.synchronize()
dt = time.perf_counter() - t
bad = 0
for i in range(steps):
    a, e, it = gather_results(outs[i]["action"], outs[i]["exitflag"], outs[i]["iters"], total)
    ga, ge, gi = got[i]
    bad += int(not (torch.equal(a, ga) and torch.equal(e, ge) and torch.equal(it, gi)))
status = x.status()
x.close()
res = torch.tensor([bad, status], device=dev)
dist.all_reduce(res, op=dist.ReduceOp.MAX)
if rank == 0:
    print(json.dumps({"world": world, "steps": steps, "depth": depth, "rows_per_rank": per, "mismatching_steps": int(res[0]), "status": int(res[1]),
                      "wall_s": dt, "ok": bool(res[0] == 0 and res[1] == 0)}))
dist.destroy_process_group()
sys.exit(0 if int(res[0]) == 0 and int(res[1]) == 0 else 1)
