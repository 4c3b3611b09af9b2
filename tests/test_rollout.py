"""Headless closed-loop rollout on the GPU: the batched planner in `previous_plan` mode drives the
MPC's own model as the plant.  Size-independent properties: nobody collides, everybody approaches the goal, warm-started
solves need fewer iterations than the cold first one, and the torch plant agrees with the solver's own prediction."""
import os

import numpy as np
import pytest
import yaml

from conftest import CONFIG, SOLVERS
from robot_mpcs_b200 import scenarios as sc

pytestmark = pytest.mark.gpu


def test_point_robot_closed_loop(built):
    import torch
    from robot_mpcs_b200.planner.mpcPlanner import MPCPlanner
    from robot_mpcs_b200.rollout import closed_loop
    from robot_mpcs_b200.utils.utils import parse_setup
    B, T = 96, 60
    setup = parse_setup(os.path.join(CONFIG, "pointRobotMpc.yaml"))["mpc"]
    setup["initialization"] = "previous_plan"
    pl = MPCPlanner("pointRobot", SOLVERS + os.sep, batch_size=B, device_resident=True, **setup)
    pl.concretize()
    pl.reset()
    pm = yaml.safe_load(open(os.path.join(SOLVERS, "pointRobot_n3_005_H20_noSlack", "paramMap.yaml")))
    s = sc.point_robot_sweep(pm, B, seed=5)
    dev = torch.device("cuda:0")
    pl._params3[:] = torch.as_tensor(s["params"], device=dev)            # scenario parameters as the set* calls leave them
    x0 = torch.as_tensor(s["xinit"], device=dev)
    P0 = s["params"][:, 0]
    goal = torch.as_tensor(P0[:, pm["goal"]], device=dev)
    obst = torch.as_tensor(P0[:, pm["obst"]], device=dev)
    rb = float(P0[0, pm["r_body"][0]])
    clear, pred_err = [], []

    def on_step(t, x, z, flag):
        # pointRobot FK: base_link sits 0.05 above the plane the obstacle centre lies in
        d = torch.sqrt(((x[:, :2] - obst[:, :2]) ** 2).sum(1) + 0.05 ** 2) - obst[:, 3] - rb
        clear.append(d.min())
        ok = flag == 1
        pred_err.append((z[ok, 1, :6] - x[ok]).abs().max())               # plant == the solver's own stage-2 state

    r = closed_loop(pl, x0, goal, T, epsilon=0.2, on_step=on_step)
    d0 = torch.linalg.norm(x0[:, :2] - goal[:, :2], dim=1)
    assert float(torch.stack(clear).min()) > -1e-6                        # no collision along any path
    assert float(torch.stack(pred_err).max()) < 1e-9
    assert int(r["failures"].sum()) == 0
    prog = (r["goal_distance"] < d0 - 0.5) | (r["goal_distance"] < 0.6) | r["reached"]
    assert float(prog.double().mean()) > 0.8                              # starts next to the goal overshoot it (1 s horizon) ...
    assert bool((r["goal_distance"] < d0 + 1.0).all())                    # ... but nobody runs away
    assert float((d0 - r["goal_distance"]).mean()) > 2.0                  # 3 s at <= 1 m/s^2: metres of progress
    it = r["mean_iters"].cpu().numpy()
    assert it[5:].mean() < 0.9 * it[0]                                    # previous-plan warm starts pay off (12.3 -> 10.2)
    assert r["path"].shape == (B, T + 1, 2)


def test_point_robot_follows_global_path(built):
    """Goal switching (boxer_example_supermarket.py:235-237): every scenario has its own 4-waypoint
    path; per step the goal parameter is the current waypoint.  Properties: the goal written into the parameters is the
    picker's waypoint, indices only move forward and by one, and scenarios get through their waypoints to the last."""
    import torch
    from robot_mpcs_b200.planner.mpcPlanner import MPCPlanner
    from robot_mpcs_b200.rollout import LocalGoalPicker, closed_loop
    from robot_mpcs_b200.utils.utils import parse_setup
    B, T, L = 64, 120, 4
    setup = parse_setup(os.path.join(CONFIG, "pointRobotMpc.yaml"))["mpc"]
    setup["initialization"] = "previous_plan"
    pl = MPCPlanner("pointRobot", SOLVERS + os.sep, batch_size=B, device_resident=True, **setup)
    pl.concretize()
    pl.reset()
    pm = yaml.safe_load(open(os.path.join(SOLVERS, "pointRobot_n3_005_H20_noSlack", "paramMap.yaml")))
    s = sc.point_robot_sweep(pm, B, seed=9)
    dev = torch.device("cuda:0")
    P = torch.as_tensor(s["params"], device=dev)
    P[:, :, pm["obst"][0]] = 50.0                                         # obstacle parked far away: this test is about goals
    P[:, :, pm["obst"][1]] = 50.0
    pl._params3[:] = P
    x0 = torch.as_tensor(s["xinit"], device=dev)
    g = torch.Generator().manual_seed(4)
    steps = torch.rand(B, L, 2, generator=g, dtype=torch.float64) * 1.5 + 0.5
    steps[..., 1] -= 1.25
    paths = (x0[:, None, :2].cpu() + torch.cumsum(steps, 1)).to(dev)
    lengths = torch.randint(2, L + 1, (B,), generator=g).to(dev)
    picker = LocalGoalPicker(paths, lengths, threshold_local_goal=1.3)
    gi = pm["goal"]
    seen = []

    def on_step(t, x, z, flag):
        ar = torch.arange(B, device=dev)
        assert torch.equal(pl._params3[:, 0, gi[:2]], paths[ar, picker.idx_local])
        assert torch.equal(pl._params3[:, -1, gi[:2]], paths[ar, picker.idx_local])
        seen.append(picker.idx_local.clone())

    r = closed_loop(pl, x0, picker.final_goal(), T, epsilon=0.3, on_step=on_step, global_path=picker)
    idx = torch.stack(seen)
    dif = idx[1:] - idx[:-1]
    assert int(dif.min()) >= 0 and int(dif.max()) <= 1
    assert bool((idx[-1] == lengths - 1).all())                           # everybody got to the last waypoint ...
    assert int(r["failures"].sum()) == 0
    d0 = torch.linalg.norm(x0[:, :2] - picker.final_goal(), dim=1)
    near = r["reached"] | (r["goal_distance"] < 2.0)                      # ... and is at it or swinging around it (the
    print("reached", float(r["reached"].double().mean()), "near", float(near.double().mean()),   # cost has no velocity term)
          "max dist", float(r["goal_distance"].max()), "d0 min", float(d0.min()))
    assert float(near.double().mean()) > 0.9
    assert bool((r["goal_distance"] < d0 + 1.5).all())


@pytest.mark.parametrize("workload,T", [("rollout_boxer_fsd", 45), ("rollout_pointRobot", 12)])
def test_bench_rollout_records(built, workload, T):
    """bench.py's closed-loop records (control steps/s, scenarios reached/s with previous_plan warm starts)."""
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import bench
    r = bench.rollout_record(workload, 128, T)
    assert r["value"] > 0 and r["failed_solves"] == 0 and r["gpu_launches"] >= 3 * T
    assert r["config"]["initialization"] == "previous_plan" and r["config"]["control_steps"] == T
    assert r["mean_iters_warm"] < r["mean_iters_first_step"]          # warm starts pay
    if workload == "rollout_boxer_fsd":
        assert r["reached_fraction"] >= 0.95 and r["scenarios_reached_per_s"] > 0


@pytest.mark.parametrize("sdir,diffdrive", [("pointRobot_n3_005_H20_noSlack", False), ("boxer_n3_01_H30_noSlack", True)])
def test_plant_kernel_is_the_models_discrete_dynamics(built, sdir, diffdrive):
    """rmpc_plant_step against the torch statement of the same model (rollout.plant_step: explicit midpoint, 5 substeps;
    models/mpcModel.py:65-69,118-120, diff_drive_mpc_model.py:24-41); failed solves are stepped with u = 0; in place."""
    import torch
    from robot_mpcs_b200.planner.solver import Solver
    from robot_mpcs_b200.rollout import plant_step
    B = 257
    solver = Solver.from_directory(os.path.join(SOLVERS, sdir), device=0, max_batch=B)
    g = torch.Generator(device="cpu").manual_seed(3)
    x = (torch.rand(B, solver.nx, generator=g, dtype=torch.float64) * 4 - 2).cuda()
    if diffdrive:
        x[:, 3:6] = 0.0
    u = (torch.rand(B, solver.nu, generator=g, dtype=torch.float64) * 2 - 1).cuda()
    flag = torch.ones(B, dtype=torch.int32, device="cuda:0")
    flag[::5] = -7
    want = plant_step(x, torch.where((flag >= 0)[:, None], u, torch.zeros_like(u)), solver.dt, diffdrive)
    got = solver.plant_step(x, u, flag)
    assert float((got - want).abs().max()) < 1e-12
    assert torch.equal(solver.plant_step(x, u), solver.plant_step(x, u, torch.ones_like(flag)))
    x2 = x.clone()
    solver.plant_step(x2, u, flag, out=x2)
    assert torch.equal(x2, got)
