#!/usr/bin/env python
"""Closed-loop rollout throughput on one B200 (the headless stand-in for examples/*_example.py): B scenarios,
T control steps each = measured state -> [per-step free-space decomposition] -> warm-started batched solve -> plant step
-> goal check, all on the device.  Prints one JSON line: control steps/s (scenario-steps), scenarios reached/s, iterations.
usage: python scripts/bench_rollout.py [--workload pointRobot|boxer_fsd] [--batch B] [--steps T]"""
import argparse, json, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch, yaml
from robot_mpcs_b200 import scenarios as sc
from robot_mpcs_b200.planner.mpcPlanner import MPCPlanner
from robot_mpcs_b200.rollout import closed_loop
from robot_mpcs_b200.utils.utils import parse_setup

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="boxer_fsd", choices=["pointRobot", "boxer_fsd"])
ap.add_argument("--batch", type=int, default=4096)
ap.add_argument("--steps", type=int, default=60)
a = ap.parse_args()
S = os.path.join(ROOT, "robot_mpcs_b200", "solvers") + os.sep
C = os.path.join(ROOT, "robot_mpcs_b200", "config")
dev = torch.device("cuda:0")
B, T = a.batch, a.steps
_stdout = sys.stdout; sys.stdout = sys.stderr            # the planner prints like the reference does


def corridor(B, seed=0):   # same scene as tests/test_gpu_features.py::_corridor_scene
    g = np.random.default_rng(seed)
    th = g.uniform(-np.pi, np.pi, B); xy = g.uniform(-1, 1, (B, 2)); L = g.uniform(4, 6, B)
    dirv = np.stack([np.cos(th), np.sin(th)], 1); nrm = np.stack([-np.sin(th), np.cos(th)], 1)
    goal = xy + L[:, None] * dirv
    along = np.linspace(-1.5, 7.5, 32)
    pts = np.concatenate([xy[:, None, :] + along[None, :, None] * dirv[:, None, :] + sgn * 1.6 * nrm[:, None, :] for sgn in (1.0, -1.0)], 1)
    pts = np.concatenate([pts, np.full((B, 64, 1), 0.02)], 2)
    xinit = np.concatenate([xy, (th + g.uniform(-0.4, 0.4, B))[:, None], np.zeros((B, 3)), g.uniform(0, 0.3, (B, 1)), np.zeros((B, 1))], 1)
    return xinit, goal, np.ascontiguousarray(pts)


def run():
    if a.workload == "pointRobot":
        setup = parse_setup(os.path.join(C, "pointRobotMpc.yaml"))["mpc"]; setup["initialization"] = "previous_plan"
        pl = MPCPlanner("pointRobot", S, batch_size=B, device_resident=True, **setup); pl.concretize(); pl.reset()
        pm = yaml.safe_load(open(S + "pointRobot_n3_005_H20_noSlack/paramMap.yaml"))
        s = sc.point_robot_sweep(pm, B, seed=5)
        pl._params3[:] = torch.as_tensor(s["params"], device=dev)
        x0 = torch.as_tensor(s["xinit"], device=dev); goal = torch.as_tensor(s["params"][:, 0, pm["goal"]], device=dev)
        return pl, x0, goal, dict(epsilon=0.2)
    setup = parse_setup(os.path.join(C, "boxerSupermarketMpc.yaml"))["mpc"]; setup["initialization"] = "previous_plan"
    xinit, goal, pts = corridor(B)
    pl = MPCPlanner("boxer", S, batch_size=B, device_resident=True, **setup); pl.concretize(); pl.reset()
    K = setup["number_obstacles"]
    pl.setLinearConstraints(np.tile([1.0, 0.0, 0.0, 100.0], (setup["time_horizon"], K, 1)), 0.6)
    pl.setJointLimits(np.array([[-30.0] * 3, [30.0] * 3])); pl.setInputLimits(np.array([[-10.0] * 2, [10.0] * 2]))
    g = torch.as_tensor(goal, device=dev); pl.setGoalReaching(g); pl.setConstraintAvoidance()
    return pl, torch.as_tensor(xinit, device=dev), g, dict(epsilon=0.4, fsd_points=torch.as_tensor(pts, device=dev))


pl, x0, goal, kw = run()
closed_loop(pl, x0, goal, 3, **kw)                      # warm-up (allocations, first launches)
pl.reset() if a.workload == "pointRobot" else None
pl, x0, goal, kw = run()
torch.cuda.synchronize()
l0 = pl._solver.launch_count()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
t = time.perf_counter(); e0.record()
r = closed_loop(pl, x0, goal, T, **kw)
e1.record(); torch.cuda.synchronize(); wall = time.perf_counter() - t
dev_s = e0.elapsed_time(e1) * 1e-3
reached = int(r["reached"].sum())
line = {"metric": "closed-loop control steps/s (scenario-steps) and scenarios reached/s", "value": B * T / dev_s, "unit": "control steps/s",
        "scenarios_reached_per_s": reached / dev_s, "reached_fraction": reached / B, "ms_per_control_step": 1e3 * dev_s / T, "wall_s": wall,
        "config": {"workload": a.workload, "batch": B, "control_steps": T, "initialization": "previous_plan",
                   "per_step": "free-space decomposition (N x K=8 planes from 64 lidar returns) + solve (N=40) + plant" if a.workload == "boxer_fsd" else "solve (N=20) + plant"},
        "mean_iters_first_step": float(r["mean_iters"][0]), "mean_iters_warm": float(r["mean_iters"][5:].mean()),
        "failed_solves": int(r["failures"].sum()), "gpu_launches": int(pl._solver.launch_count() - l0), "dtype": "f64", "data": "synthetic"}
print(json.dumps(line), file=_stdout)
