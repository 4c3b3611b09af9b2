"""MPCPlanner — drop-in for robotmpcs.planner.mpcPlanner.MPCPlanner (reference planner/mpcPlanner.py:31-301) whose
solver call runs B independent problems on the GPU.

Same constructor, same `reset/concretize/set*/solve/computeAction` methods and argument meanings, same solver
directory contract (`<solversDir><robotType>_n<n>_<dt>_H<N>[_noSlack]/{paramMap.yaml,properties.yaml}`), same
error behaviour (SolverDoesNotExistError when the directory is missing; a negative exitflag is printed and
returned, never raised).  With `batch_size == 1` and numpy inputs every return value has the reference's shape
(`action[nu]`, `output{'x01'..}`, `exitflag:int`).  With `batch_size > 1` every `set*` takes a leading batch
dimension (or broadcasts a single scenario), and `solve/computeAction` return `action[B,nu]`, `z[B,N,nvar]`,
`exitflag[B]`.  With `device_resident=True` the parameter / warm-start buffers are CUDA tensors and nothing
crosses PCIe between control steps except what the caller passes in.

The reference's defects listed in SURVEY.md §3.6 are not reproduced (EmptyObstacle.dimension, `_r`, the
3-tuple/2-tuple mismatch is kept as the 3-tuple the method actually returns).
"""
from __future__ import annotations

import os
from typing import Tuple

import numpy as np
import yaml

from robot_mpcs_b200.models.mpcBase import MpcConfiguration
from robot_mpcs_b200.planner.solver import Solver

try:
    import torch
except Exception:  # pragma: no cover
    torch = None


class SolverDoesNotExistError(Exception):  # mpcPlanner.py:10-16
    def __init__(self, solverName):
        super().__init__()
        self._solverName = solverName

    def __str__(self):
        return f"Solver with name {self._solverName} does not exist."


class EmptyObstacle:  # mpcPlanner.py:18-26: padding obstacle, constraint trivially inactive
    def position(self):
        return [-100, -100, -100]

    def radius(self):
        return -100

    def dimension(self):
        return 3

    dim = dimension


class PlannerSettingIncomplete(Exception):
    pass


class MPCPlanner(object):
    def __init__(self, robotType, solversDir, mpc_model=None, debug=False, batch_size=1, device=0,
                 device_resident=False, **kwargs):
        self._config = MpcConfiguration(**kwargs)
        self._robotType = robotType
        self._initial_step = True
        self._debug = debug
        self._B = int(batch_size)
        self._device_resident = bool(device_resident)
        self._device = device
        dt_str = str(self._config.time_step).replace(".", "")
        self._solverFile = (solversDir + self._robotType + "_n" + str(self._config.n) + "_" + dt_str + "_H"
                            + str(self._config.time_horizon))
        if not self._config.slack:
            self._solverFile += "_noSlack"
        if not os.path.isdir(self._solverFile):
            raise SolverDoesNotExistError(self._solverFile)
        with open(self._solverFile + "/paramMap.yaml", "r") as stream:
            self._paramMap = yaml.safe_load(stream)
        with open(self._solverFile + "/properties.yaml", "r") as stream:
            self._properties = yaml.safe_load(stream)
        self._nx = self._properties["nx"]
        self._nu = self._properties["nu"]
        self._ns = self._properties["ns"]
        self._npar = self._properties["npar"]
        try:
            print("Loading solver %s" % self._solverFile)
            self._solver = Solver.from_directory(self._solverFile, device=device, max_batch=self._B)
        except Exception as e:
            print("FAILED TO LOAD SOLVER")
            raise e
        self._solver.set_options(control_mode=1 if self._config.control_mode == "vel" else 0)
        if self._debug:
            self._mpc_model = mpc_model
        self._r = 0.1
        self._actionCounter = self._config.interval
        self._z_prev = None
        self.output = None

    # ---- buffers ---------------------------------------------------------------------------
    def _zeros(self, *shape):
        if self._device_resident:
            return torch.zeros(*shape, dtype=torch.float64, device=f"cuda:{self._device}")
        return np.zeros(shape, dtype=np.float64)

    def _as(self, v, ndim_single):
        """Bring a user value to the buffer's array type with a leading batch dimension (broadcast if absent)."""
        if self._device_resident:
            t = torch.as_tensor(np.asarray(v, dtype=np.float64) if not isinstance(v, torch.Tensor) else v,
                                dtype=torch.float64, device=f"cuda:{self._device}")
        else:
            t = np.asarray(v.cpu() if (torch is not None and isinstance(v, torch.Tensor)) else v, dtype=np.float64)
        if t.ndim == ndim_single:
            t = t[None]
        return t

    def _set(self, name, values, stage_varying=False):
        """params[:, :, paramMap[name]] = values ([B|1, len] or [B|1, N, len]).  Device resident: one launch of the
        library's scatter kernel (rmpc_set_params) — the set* loops of mpcPlanner.py:120-210 for the whole batch."""
        idx = self._paramMap[name]
        v = self._as(values, 2 if stage_varying else 1)
        if self._device_resident:
            shape = (self._B, self._config.time_horizon, len(idx)) if stage_varying else (self._B, len(idx))
            self._solver.set_params(self._params3, idx, v.expand(*shape))
            return
        if not stage_varying:
            v = v[:, None, :]
        self._params3[:, :, idx] = v

    @property
    def _params(self):
        """Reference-shaped flat view for B == 1 (mpcPlanner.py:91); [B, N*npar] otherwise."""
        p = self._params3.reshape(self._B, -1)
        return p[0] if self._B == 1 else p

    def reset(self):  # mpcPlanner.py:83-104
        print("RESETTING PLANNER")
        N = self._config.time_horizon
        self._x0 = self._zeros(self._B, N, self._nx + self._nu + self._ns)
        self._xinit = self._zeros(self._B, self._nx)
        self._initial_step = True
        self._params3 = self._zeros(self._B, N, self._npar)
        w = self._config.weights
        if "wgoal" in self._paramMap:
            self._set("wgoal", np.ones(len(self._paramMap["wgoal"])) * w["w"])  # weights['w'], as the reference (:93-95)
        self._set("wu", np.ones(len(self._paramMap["wu"])) * np.asarray(w["wu"], dtype=np.float64))
        self._z_prev = None

    def m(self):
        return self._properties["m"]

    def dt(self):
        return self._config.time_step

    # ---- set* (mpcPlanner.py:120-210) ----------------------------------------------------------
    def setRadialConstraints(self, obsts, r_body):
        """obsts: list of objects with position()/radius() (one scenario, broadcast), or an array
        [B, n_obst, 4] of (x, y, z, r); missing obstacles are padded with EmptyObstacle."""
        K = self._config.number_obstacles
        self._set("r_body", np.atleast_1d(r_body) if np.ndim(r_body) == 0 else np.asarray(r_body).reshape(-1, 1))
        if isinstance(obsts, (list, tuple)) and (len(obsts) == 0 or hasattr(obsts[0], "position")):
            arr = np.zeros((1, K, 4))
            for j in range(K):
                ob = obsts[j] if j < len(obsts) else EmptyObstacle()
                pos = list(ob.position())
                arr[0, j, :3] = (pos + [0.0, 0.0, 0.0])[:3]
                arr[0, j, 3] = ob.radius()
        else:
            arr = self._as(obsts, 2)
            if arr.shape[1] < K:
                pad = self._zeros(arr.shape[0], K - arr.shape[1], 4) - 100.0
                arr = (torch.cat if self._device_resident else np.concatenate)([arr, pad], 1)
        self._set("obst", arr.reshape(arr.shape[0], 4 * K))

    def setObstacles(self, obsts, r_body):  # legacy name still used by the ROS node (mpc_planner_node:80)
        self.setRadialConstraints(obsts, r_body)

    def setLinearConstraints(self, lin_constr, r_body):
        """lin_constr[N][K][4] (one scenario) or [B][N][K][4]; stage-varying (mpcPlanner.py:135-141)."""
        self._set("r_body", np.atleast_1d(r_body) if np.ndim(r_body) == 0 else np.asarray(r_body).reshape(-1, 1))
        lc = self._as(lin_constr, 3)
        for i in range(self._config.number_obstacles):
            self._set("lin_constrs_" + str(i), lc[:, :, i, :], stage_varying=True)

    def updateDynamicObstacles(self, obstArray):
        """obstArray[(B,) 3*m*n_dyn] = per obstacle (pos, vel, acc); stage i sees pos + vel (i dt) + acc (i dt)^2 / 2 with
        radius self._r, obstacle slots beyond the moving ones are parked at -100 (mpcPlanner.py:144-161; the reference
        reads the first obstacle's block for every j and calls a missing self.dt() — its evident intent is followed).
        Device resident: rmpc_predict_obstacles, one kernel for the whole batch."""
        a = self._as(obstArray, 1)
        m, K, N = self.m(), self._config.number_obstacles, self._config.time_horizon
        nb = min(K, a.shape[1] // (3 * m))
        p_obst = self._paramMap["obst"][0]
        if self._device_resident:
            st = a[:, : nb * 3 * m].reshape(a.shape[0], nb, 3 * m).expand(self._B, nb, 3 * m) if nb else None
            self._solver.predict_obstacles(self._params3, st, K, p_obst, self._r)
            return
        i = np.arange(N, dtype=np.float64)[None, :, None]
        for j in range(K):
            idx = self._paramMap["obst"][j * (m + 1): (j + 1) * (m + 1)]
            if j < nb:
                blk = a[:, j * 3 * m: (j + 1) * 3 * m]
                pos, vel, acc = blk[:, None, :m], blk[:, None, m:2 * m], blk[:, None, 2 * m:3 * m]
                pred = pos + vel * self.dt() * i + 0.5 * (self.dt() * i) ** 2 * acc       # mpcPlanner.py:158, same operation order
            else:
                pred = np.zeros((1, N, m)) - 100.0
            self._params3[:, :, idx[:m]] = pred
            self._params3[:, :, idx[m]] = self._r

    def setSelfCollisionAvoidanceConstraints(self, r_body):
        self._set("r_body", np.atleast_1d(r_body) if np.ndim(r_body) == 0 else np.asarray(r_body).reshape(-1, 1))

    def setJointLimits(self, limits):
        lim = self._as(limits, 2)
        self._set("lower_limits", lim[:, 0, : self._config.n])
        self._set("upper_limits", lim[:, 1, : self._config.n])

    def setVelLimits(self, limits_vel):
        lim = self._as(limits_vel, 2)
        self._set("lower_limits_vel", lim[:, 0, :2])
        self._set("upper_limits_vel", lim[:, 1, :2])

    def setInputLimits(self, limits_u):
        lim = self._as(limits_u, 2)
        self._set("lower_limits_u", lim[:, 0, : self._nu])
        self._set("upper_limits_u", lim[:, 1, : self._nu])

    def setGoalReaching(self, goal_position):
        g = self._as(goal_position, 1)
        full = self._zeros(g.shape[0], self.m())
        k = min(g.shape[1], self.m())
        full[:, :k] = g[:, :k]  # missing dimensions padded with 0 (mpcPlanner.py:199-204)
        self._set("goal", full)

    def setGoal(self, goal):  # legacy: GoalComposition (mpc_planner_node:116)
        self.setGoalReaching(goal.primary_goal().position())

    def setConstraintAvoidance(self):
        self._set("wconstr", np.asarray(self._config.weights["wconstr"], dtype=np.float64))

    # ---- warm start (mpcPlanner.py:215-236) ----------------------------------------------------
    def shiftHorizon(self, output):
        z = output
        if isinstance(output, dict):
            z = np.stack([output[k] for k in sorted(output.keys())])[None]
        if self._device_resident:
            self._solver.shift_horizon(z, self._x0)
        else:
            self._x0[:, :-1, :] = z[:, 1:, :]
            self._x0[:, -1, :] = z[:, -1, :]

    def setX0(self, initialize_type="current_state", initial_step=True):
        if initialize_type == "current_state" or (initialize_type == "previous_plan" and initial_step):
            if self._device_resident:
                self._solver.set_x0(self._xinit, self._x0)
            else:
                self._x0[:, :, : self._nx] = self._xinit[:, None, :]
            self._initial_step = False
        elif initialize_type == "previous_plan":
            self.shiftHorizon(self._z_prev)

    # ---- solve -----------------------------------------------------------------------------------
    def solve(self, ob):
        ob = self._as(ob, 1)
        if ob.shape[0] != self._B:
            raise ValueError(f"observation batch {ob.shape[0]} != batch_size {self._B}")
        self._xinit = ob[:, : self._nx]
        if ob.shape[1] > self._nx:
            self.updateDynamicObstacles(ob[:, self._nx:])
        self.setX0(initialize_type=self._config.initialization, initial_step=self._initial_step)
        res = self._solver.solve_batch(self._xinit, self._x0, self._params3)
        self._z_prev = res["z"]
        exitflag = res["exitflag"]
        info = {k: res[k] for k in ("iters", "obj", "res_stat", "res_eq", "res_ineq", "res_comp")}
        if self._config.control_mode not in ("vel", "acc"):
            print("No valid control mode specified!")
            action = self._zeros(self._B, self._nu)
        else:
            action = res["action"]
        if self._B == 1 and not self._device_resident:
            flag = int(exitflag[0])
            if flag < 0:
                print(flag)
            width = len(str(self._config.time_horizon))
            self.output = {"x" + str(k + 1).zfill(width): res["z"][0, k] for k in range(self._config.time_horizon)}
            return action[0], self.output, {k: v[0] for k, v in info.items()}, flag
        self.output = res["z"]
        return action, self.output, info, exitflag

    def concretize(self):  # mpcPlanner.py:290-291
        self._actionCounter = self._config.interval

    def computeAction(self, *args) -> Tuple[np.ndarray, dict, int]:
        """computeAction(q, qdot[, qudot]) -> (action, output, exitflag)  (mpcPlanner.py:293-301)."""
        cat = torch.cat if (torch is not None and isinstance(args[0], torch.Tensor)) else np.concatenate
        ob = cat(args[:3], -1)
        if self._actionCounter >= self._config.interval:
            self._action, self._last_output, _info, self._last_flag = self.solve(ob)
            self._last_iters = _info["iters"]
            self._actionCounter = 1
        else:
            self._actionCounter += 1
        return self._action, self._last_output, self._last_flag
