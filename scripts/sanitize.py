"""Small solves of every model through the C ABI, for compute-sanitizer (memcheck / racecheck / synccheck / initcheck):
  compute-sanitizer --tool racecheck python scripts/sanitize.py
Covers both solve kernels (feasible and infeasible starts), the compressed parameter format, the warm-start shift, the
parameter scatter and the free-space decomposition."""
import os, sys, yaml, numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from robot_mpcs_b200 import scenarios as sc
from robot_mpcs_b200.planner.solver import Solver
S = os.path.join(ROOT, "robot_mpcs_b200", "solvers")
CASES = [("pointRobot_n3_005_H20_noSlack", sc.point_robot_sweep, 16), ("boxer_n3_01_H30_noSlack", sc.boxer_radial, 16),
         ("panda_n7_005_H20_noSlack", sc.panda, 8), ("boxer_n3_01_H40_noSlack", sc.boxer_supermarket, 8)]
for sdir, gen, B in CASES:
    pm = yaml.safe_load(open(f"{S}/{sdir}/paramMap.yaml"))
    s = gen(pm, B)
    solver = Solver.from_directory(f"{S}/{sdir}", device=0, max_batch=B)
    solver.set_options(max_iter=40)
    d = {k: torch.as_tensor(np.ascontiguousarray(v), device="cuda:0") for k, v in s.items() if isinstance(v, np.ndarray)}
    r = solver.solve_batch(d["xinit"], d["x0"], d["params"])
    x0 = solver.shift_horizon(r["z"])
    # the shifted plan against displaced obstacles / planes: some rows start violated -> the kernel for infeasible starts
    p2 = d["params"].clone()
    if "obst" in pm:
        p2[:, :, pm["obst"][0]:pm["obst"][0] + 2] = r["z"][:, 10:11, 0:2]
    r2 = solver.solve_batch(d["xinit"], x0, p2)
    if bool((d["params"] == d["params"][:, :1]).all()):
        r3 = solver.solve_batch(d["xinit"], d["x0"], d["params"][:, 0].contiguous())
        assert torch.equal(r3["z"], r["z"])
    if sdir.endswith("H40_noSlack"):
        pts, ref = sc.lidar_scene(B, solver.N, 64, seed=1)
        idx = [pm[f"lin_constrs_{i}"][0] for i in range(8)]
        solver.fsd_constraints(torch.as_tensor(pts, device="cuda:0"), torch.as_tensor(ref, device="cuda:0"), 8, 5.0, params=p2, idx=idx)
        solver.solve_batch(d["xinit"], x0, p2)
    h = solver.solve_batch(s["xinit"], s["x0"], s["params"])           # host entry point
    torch.cuda.synchronize()
    print(sdir, "flags", np.unique(r["exitflag"].cpu().numpy(), return_counts=True), "after displacement", np.unique(r2["exitflag"].cpu().numpy(), return_counts=True))
print("done")
