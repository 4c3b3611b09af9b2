"""Headless, device-resident closed-loop rollout of B scenarios: the stand-in for the reference's
pybullet demos (examples/pointRobot_example.py:81-89, boxer_example.py:176-224 — the simulator and its GUI are
unavailable on this box).  Per control step and for the whole batch: measured state -> MPCPlanner.computeAction (warm
start, batched solve on the GPU) -> the first action drives the plant -> goal check.  The plant is the MPC's own
discrete model (double integrator, or diff-drive with explicit-midpoint substeps), integrated here in torch on the
device; nothing leaves the GPU until the summary at the end.

    reference                                                   here
    env.step(action) (pybullet)                                  plant_step(x, u)            (model = plant)
    check_goal_reaching, boxer_example.py:219-224                ||q[:2] - goal|| <= epsilon, per scenario
    per-stage FreeSpaceDecomposition in Python, :191-203         fsd_points=... -> rmpc_fsd_constraints on the previous plan
    GlobalPlanner.get_local_goal + setGoalReaching,              global_path=LocalGoalPicker(...): waypoint index per
      boxer_example_supermarket.py:235-237                         scenario kept on the device, goal parameter rewritten per step
"""
from __future__ import annotations

import math

import torch


class LocalGoalPicker:
    """Batched GlobalPlanner.get_local_goal (robotmpcs/global_planner/globalPlanner.py:174-189, threshold :13): every
    scenario walks along its own global path; the waypoint index moves on by one when the robot is within
    `threshold_local_goal` of the current waypoint and the waypoint is not the last.  The A* search that produces the
    path is out of scope; paths come in as data.

    paths[B, L, 2] (rows beyond a scenario's length are ignored), lengths[B] >= 1; tensors live wherever `paths` does.
    With B == 1 and a python list/ndarray path, `get_local_goal(position, path)` has the reference's signature."""

    def __init__(self, paths=None, lengths=None, threshold_local_goal: float = 1.3):
        self.threshold_local_goal = float(threshold_local_goal)
        self.idx_local = 0
        self._paths = None
        if paths is not None:
            self.set_paths(paths, lengths)

    def set_paths(self, paths, lengths=None):
        paths = torch.as_tensor(paths, dtype=torch.float64)
        if paths.ndim == 2:
            paths = paths[None]
        B, L = paths.shape[0], paths.shape[1]
        if lengths is None:
            lengths = torch.full((B,), L, dtype=torch.int64, device=paths.device)
        lengths = torch.as_tensor(lengths, dtype=torch.int64, device=paths.device)
        if L == 0 or bool((lengths < 1).any()) or bool((lengths > L).any()):
            raise IndexError("every scenario needs a path of 1..L waypoints")  # the reference indexes path[0]: IndexError
        self._paths, self._lengths = paths[..., :2].contiguous(), lengths
        self.idx_local = torch.zeros(B, dtype=torch.int64, device=paths.device)
        return self

    def step(self, position: torch.Tensor) -> torch.Tensor:
        """position[B, >=2] -> local_goal[B, 2]; advances idx_local in place."""
        B = self._paths.shape[0]
        ar = torch.arange(B, device=self._paths.device)
        wp = self._paths[ar, self.idx_local]
        d = torch.sqrt((wp[:, 0] - position[:, 0]) ** 2 + (wp[:, 1] - position[:, 1]) ** 2)
        adv = (self.idx_local < self._lengths - 1) & (d <= self.threshold_local_goal)
        self.idx_local = self.idx_local + adv.to(torch.int64)
        return self._paths[ar, self.idx_local]

    def final_goal(self) -> torch.Tensor:
        ar = torch.arange(self._paths.shape[0], device=self._paths.device)
        return self._paths[ar, self._lengths - 1]

    def get_local_goal(self, position, path):
        """Reference-shaped single-scenario call: the path is passed every time, the index is kept by the object."""
        i = int(self.idx_local[0]) if isinstance(self.idx_local, torch.Tensor) else int(self.idx_local)
        wp = path[i]                                     # empty path: IndexError, as in the reference (:181)
        d = math.sqrt((wp[0] - position[0]) ** 2 + (wp[1] - position[1]) ** 2)
        if i < len(path) - 1 and d <= self.threshold_local_goal:
            i += 1
        self.idx_local = i
        return path[i]


def plant_step(x: torch.Tensor, u: torch.Tensor, dt: float, diffdrive: bool, nodes: int = 5) -> torch.Tensor:
    """x+ of the model the solver was generated from (robotmpcs/models/mpcModel.py:65-69,118-120 and
    diff_drive_mpc_model.py:24-41): ERK2 = explicit midpoint, `nodes` substeps."""
    if not diffdrive:
        n = x.shape[1] // 2
        q, qd = x[:, :n], x[:, n:]
        return torch.cat([q + dt * qd + 0.5 * dt * dt * u, qd + dt * u], 1)
    h = dt / nodes
    x = x.clone()
    for _ in range(nodes):
        def f(s):
            th, v, om = s[:, 2], s[:, 6], s[:, 7]
            z = torch.zeros_like(th)
            return torch.stack([torch.cos(th) * v, torch.sin(th) * v, om, z, z, z, u[:, 0], u[:, 1]], 1)
        k1 = f(x)
        x = x + h * f(x + 0.5 * h * k1)
    return x


def closed_loop(planner, x0: torch.Tensor, goal: torch.Tensor, steps: int, epsilon: float = 0.2, fsd_points=None,
                fsd_radius: float = 5.0, on_step=None, global_path: "LocalGoalPicker | None" = None) -> dict:
    """Runs `steps` control steps for the planner's whole batch.  x0[B,nx], goal[B,>=2] CUDA float64; the planner must
    be device resident with its scenario parameters already set.  fsd_points[B,P,3] switches on the per-step
    free-space decomposition (LinearConstraints models): planes for stage k are seeded from the previous plan's stage
    k+1 pose, or from the measured pose at the first step, as boxer_example.py:191-201 does.  global_path: a
    LocalGoalPicker on the same device; per step the goal parameter becomes the scenario's current waypoint
    (boxer_example_supermarket.py:235-237) and `goal` (the final goal) is only used for the reached check."""
    diffdrive = planner._robotType.startswith("boxer")
    dt = planner.dt()
    nx, B = planner._nx, x0.shape[0]
    n = planner._config.n
    x = x0.clone()
    reached = torch.zeros(B, dtype=torch.bool, device=x.device)
    first_reached = torch.full((B,), -1, dtype=torch.int64, device=x.device)
    iters = torch.zeros(steps, dtype=torch.float64, device=x.device)
    failures = torch.zeros(B, dtype=torch.int64, device=x.device)
    traj = [x[:, :2].clone()]
    K = planner._config.number_obstacles
    idx = [planner._paramMap[f"lin_constrs_{i}"][0] for i in range(K)] if fsd_points is not None else None
    z_prev = None
    for t in range(steps):
        if global_path is not None:
            planner.setGoalReaching(global_path.step(x[:, :2]))
        if fsd_points is not None:
            if z_prev is None:
                ref = x[:, :3].contiguous()
            else:  # stage k of the new problem ~ stage k+1 of the previous plan
                ref = planner._solver.shift_horizon(z_prev)
            planner._solver.fsd_constraints(fsd_points, ref, K, fsd_radius, params=planner._params3, idx=idx, want_out=False)
        if diffdrive:
            action, z, flag = planner.computeAction(x[:, :n], x[:, n:2 * n], x[:, 2 * n:])
        else:
            action, z, flag = planner.computeAction(x[:, :n], x[:, n:])
        info_iters = planner._last_iters if hasattr(planner, "_last_iters") else None
        if info_iters is not None:
            iters[t] = info_iters.to(torch.float64).mean()
        if planner._config.control_mode == "vel":
            # the action is the velocity planned for stage 2 (mpcPlanner.py:275-278); the plant is driven by inputs, and for
            # these models v+ = v + dt u, so the input that realises it is (action - v) / dt — the plan's own u_1
            vel = x[:, 2 * n:2 * n + action.shape[1]] if diffdrive else x[:, n:2 * n]
            action = (action - vel) / dt
        failures += (flag < 0).to(torch.int64)
        # the plant is the model the solver was generated from, stepped by the library's own kernel (one launch instead of the
        # ~100 elementwise launches of `plant_step` above, which stays as the torch statement of the same model for the tests);
        # boxer_example.py:194: a failed solve is not applied
        x = planner._solver.plant_step(x, action, flag)
        z_prev = z
        d = torch.linalg.norm(x[:, :2] - goal[:, :2], dim=1)
        newly = (d <= epsilon) & ~reached
        first_reached = torch.where(newly, torch.full_like(first_reached, t + 1), first_reached)
        reached |= newly
        traj.append(x[:, :2].clone())
        if on_step is not None:
            on_step(t, x, z, flag)
    return {"x": x, "reached": reached, "first_reached": first_reached, "mean_iters": iters, "failures": failures,
            "goal_distance": torch.linalg.norm(x[:, :2] - goal[:, :2], dim=1), "path": torch.stack(traj, 1)}
