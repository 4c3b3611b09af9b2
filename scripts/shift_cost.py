"""What the problems with an infeasible initial guess cost a batch: panda, B = 1024, one stream; the batch as generated
(7 such problems) against the same batch with those problems replaced by copies of a feasible one."""
import os, sys, yaml, numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from robot_mpcs_b200 import scenarios as sc
from robot_mpcs_b200.planner.solver import Solver
S = os.path.join(ROOT, "robot_mpcs_b200", "solvers")
sdir, B = "panda_n7_005_H20_noSlack", 1024
pm = yaml.safe_load(open(f"{S}/{sdir}/paramMap.yaml"))
s = sc.panda(pm, B)
solver = Solver.from_directory(f"{S}/{sdir}", device=0, max_batch=B)
d = {k: torch.as_tensor(v, device="cuda:0") for k, v in s.items() if isinstance(v, np.ndarray)}
def run(d, n=20):
    for _ in range(3): r = solver.solve_batch(d["xinit"], d["x0"], d["params"])
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): r = solver.solve_batch(d["xinit"], d["x0"], d["params"])
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n, r
t, r = run(d)
ef = r["exitflag"].cpu().numpy(); it = r["iters"].cpu().numpy()
print("as generated: %.3f ms/batch" % t, "flags", dict(zip(*np.unique(ef, return_counts=True))), "max iters", it.max(), "iters of failed", it[ef != 1])
bad = np.where(ef != 1)[0]
slow = np.argsort(-it)[:12]
print("slowest", slow, it[slow], ef[slow])
d2 = {k: torch.as_tensor(np.ascontiguousarray(v.cpu().numpy()).copy(), device="cuda:0") for k, v in d.items()}
good = int(np.where((ef == 1) & (it <= 12))[0][0])
for k in d2:
    d2[k][torch.as_tensor(bad, device="cuda:0")] = d2[k][good].clone()
t2, r2 = run(d2)
print("failed problems replaced by a feasible one: %.3f ms/batch" % t2, "max iters", int(r2["iters"].max()))
