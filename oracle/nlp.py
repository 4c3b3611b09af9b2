"""ORACLE — test infrastructure only (never imported by the product path).

Float64 CPU restatement of the NLP that robotmpcs builds symbolically with CasADi.  Written with
torch.float64 on the CPU so that exact first derivatives come from autograd (the role CasADi AD plays
in the reference).

PINNED to outputs of the reference itself, run here: tests/golden/model_reference.json holds what the reference's own
model classes return for inequalities, stage / terminal cost, continuous dynamics, paramMap, dimensions, bounds and
solver name (tests/golden/make_model_golden.py: CasADi replaced by the same arithmetic on numbers);
tests/test_model_reference.py checks this file against them to rounding.
NOT PINNED (third-party code absent here and on the GPU box): the URDF forward kinematics
(forwardkinematics 1.1.5 — ChainFk below restates the URDF convention; FK anchors of SURVEY.md §3.5 in
tests/test_oracle_model.py), the ERK2 tableau and the solve itself (FORCESPRO; see solve.py).

Each function cites the reference file:line it restates (paths relative to /root/reference).
"""
from __future__ import annotations

import math
import xml.etree.ElementTree as ET
from dataclasses import dataclass, field
from typing import Dict, List, Tuple

import numpy as np
import torch
import yaml

DT = torch.float64


# ----------------------------------------------------------------------------------------------
# URDF chain forward kinematics — stands in for forwardkinematics.GenericURDFFk (3rd party, v1.1.5,
# absent).  Call sites restated: models/mpcBase.py:46-51,89-94, objectives/goal_reaching.py:22-27.
# ----------------------------------------------------------------------------------------------
@dataclass
class _Joint:
    name: str
    jtype: str
    parent: str
    child: str
    xyz: Tuple[float, float, float]
    rpy: Tuple[float, float, float]
    axis: Tuple[float, float, float]


def _rpy_matrix(r, p, y):
    # URDF fixed-axis roll/pitch/yaw: R = Rz(y) Ry(p) Rx(r)
    cr, sr, cp, sp, cy, sy = math.cos(r), math.sin(r), math.cos(p), math.sin(p), math.cos(y), math.sin(y)
    return np.array([
        [cy * cp, cy * sp * sr - sy * cr, cy * sp * cr + sy * sr],
        [sy * cp, sy * sp * sr + cy * cr, sy * sp * cr - cy * sr],
        [-sp, cp * sr, cp * cr],
    ])


class ChainFk:
    """fk(q, root, link, positionOnly=True) for a serial/tree URDF.  q is indexed by the movable
    joints met on the root->end chain, in order (that count is `n()`)."""

    def __init__(self, urdf_text: str, root_link: str, end_link: str, base_type: str = "holonomic"):
        robot = ET.fromstring(urdf_text)
        self.joints: Dict[str, _Joint] = {}
        self.parent_joint: Dict[str, _Joint] = {}
        for j in robot.findall("joint"):
            org = j.find("origin")
            xyz = tuple(float(v) for v in (org.get("xyz", "0 0 0") if org is not None else "0 0 0").split())
            rpy = tuple(float(v) for v in (org.get("rpy", "0 0 0") if org is not None else "0 0 0").split())
            ax = j.find("axis")
            axis = tuple(float(v) for v in (ax.get("xyz") if ax is not None else "1 0 0").split())
            jt = _Joint(j.get("name"), j.get("type"), j.find("parent").get("link"), j.find("child").get("link"), xyz, rpy, axis)
            self.joints[jt.name] = jt
            self.parent_joint[jt.child] = jt
        self.links = [l.get("name") for l in robot.findall("link")]
        self.root = root_link
        self.end = end_link
        self.base_type = base_type
        self.q_index: Dict[str, int] = {}
        for jt in self._chain(root_link, end_link):
            if jt.jtype != "fixed":
                self.q_index[jt.name] = len(self.q_index)

    def _chain(self, root: str, link: str) -> List[_Joint]:
        out = []
        cur = link
        while cur != root:
            if cur not in self.parent_joint:
                raise KeyError(f"link {link} is not below {root}")
            jt = self.parent_joint[cur]
            out.append(jt)
            cur = jt.parent
        return out[::-1]

    def n(self) -> int:
        return len(self.q_index)

    def fk(self, q: torch.Tensor, root: str, link: str) -> torch.Tensor:
        """q[..., n_total] -> position [..., 3] of `link` expressed in `root` (pre-multiplied by the
        planar base pose for diffdrive, whose first three coordinates are X, Y, theta)."""
        lead = q.shape[:-1]
        R = torch.eye(3, dtype=DT).expand(*lead, 3, 3)
        t = torch.zeros(*lead, 3, dtype=DT)
        off = 0
        if self.base_type == "diffdrive":
            c, s = torch.cos(q[..., 2]), torch.sin(q[..., 2])
            z0, o1 = torch.zeros_like(c), torch.ones_like(c)
            R = torch.stack([torch.stack([c, -s, z0], -1), torch.stack([s, c, z0], -1), torch.stack([z0, z0, o1], -1)], -2)
            t = torch.stack([q[..., 0], q[..., 1], z0], -1)
            off = 3
        for jt in self._chain(root, link):
            Ro = torch.tensor(_rpy_matrix(*jt.rpy), dtype=DT)
            to = torch.tensor(jt.xyz, dtype=DT)
            t = t + (R @ to)
            R = R @ Ro
            if jt.jtype == "fixed":
                continue
            qi = q[..., off + self.q_index[jt.name]]
            a = torch.tensor(jt.axis, dtype=DT)
            a = a / a.norm()
            if jt.jtype == "prismatic":
                t = t + (R @ a) * qi[..., None]
            else:  # revolute / continuous: Rodrigues
                K = torch.tensor([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]], dtype=DT)
                Rj = torch.eye(3, dtype=DT) + torch.sin(qi)[..., None, None] * K + (1 - torch.cos(qi))[..., None, None] * (K @ K)
                R = R @ Rj
        return t


# ----------------------------------------------------------------------------------------------
# Model = dims + paramMap + J, h, Phi  (models/mpcBase.py, mpcModel.py, diff_drive_mpc_model.py,
# inequalities/*, objectives/*)
# ----------------------------------------------------------------------------------------------
class OracleModel:
    def __init__(self, setup: dict, urdf_text: str):
        mpc, rob = setup["mpc"], setup["robot"]
        self.mpc, self.rob = mpc, rob
        self.fkm = ChainFk(urdf_text, rob["root_link"], rob["end_link"], rob["base_type"])
        self.N = int(mpc["time_horizon"])
        self.dt = float(mpc["time_step"])
        self.m = 3
        # models/mpcBase.py:52-66
        if rob["base_type"] == "holonomic":
            self.n = self.fkm.n()
            self.nx = 2 * self.n
            self.nu = self.n
        else:
            self.n = self.fkm.n() + 3
            self.nx = 2 * self.n + 2
            self.nu = 2 + self.fkm.n()
        self.ns = 0
        self.nvar = self.nx + self.nu
        self.K = int(mpc["number_obstacles"])
        self.links = list(rob["collision_links"])
        self.pairs = [tuple(p) for p in rob["selfCollision"]["pairs"]]
        self.constraints = list(mpc["constraints"])
        self.objectives = list(mpc["objectives"])
        # paramMap: first-come allocation, duplicates skipped (models/mpcBase.py:68-71); order =
        # constraints (mpcModel.py:29-30) -> wu (ObjectiveManager.py:14) -> objectives (mpcModel.py:35-36)
        self.paramMap: Dict[str, List[int]] = {}
        self.npar = 0
        for c in self.constraints:
            if c == "RadialConstraints":            # inequalities/RadialConstraints.py:15-16
                self._add("r_body", 1); self._add("obst", 4 * self.K)
            elif c == "LinearConstraints":          # inequalities/LinearConstraints.py:21-24
                self._add("r_body", 1)
                for i in range(self.K):
                    self._add(f"lin_constrs_{i}", 4)
            elif c == "SelfCollisionAvoidanceConstraints":  # :16
                self._add("r_body", 1)
            elif c == "JointLimitConstraints":      # inequalities/JointLimitConstraints.py:15-16
                self._add("lower_limits", self.n); self._add("upper_limits", self.n)
            elif c == "InputLimitConstraints":      # inequalities/InputLimitConstraints.py:13-14
                self._add("lower_limits_u", self.nu); self._add("upper_limits_u", self.nu)
            elif c == "VelLimitConstraints":        # inequalities/VelLimitConstraints.py:14-15
                self._add("lower_limits_vel", 2); self._add("upper_limits_vel", 2)
            else:
                raise ValueError(c)
        self._add("wu", self.nu)
        for o in self.objectives:
            if o == "GoalReaching":                 # objectives/goal_reaching.py:14-15
                self._add("goal", self.m); self._add("wgoal", self.m)
            elif o == "ConstraintAvoidance":        # objectives/constraint_avoidance.py:22
                self._add("wconstr", len(self.constraints))
            else:
                raise ValueError(o)
        self.nh = int(self.ineq(torch.zeros(1, self.nvar, dtype=DT) + 0.1, torch.ones(1, self.npar, dtype=DT)).shape[-1])

    def _add(self, name, k):
        if name not in self.paramMap:
            self.paramMap[name] = list(range(self.npar, self.npar + k))
            self.npar += k

    def solver_name(self) -> str:  # models/mpcModel.py:111-113
        s = f"{self.mpc['model_name']}_n{self.n}_{str(self.dt).replace('.', '')}_H{self.N}"
        return s + ("" if self.mpc["slack"] else "_noSlack")

    # ---- pieces ------------------------------------------------------------------------------
    def _P(self, p, name):
        return p[..., self.paramMap[name]]

    def _fk(self, z, link):
        return self.fkm.fk(z[..., : self.n], self.rob["root_link"], link)

    def _module_rows(self, name, z, p) -> List[torch.Tensor]:
        rows = []
        if name == "RadialConstraints":  # models/mpcBase.py:82-101 (intent of RadialConstraints.py:21-23)
            ob, rb = self._P(p, "obst"), self._P(p, "r_body")[..., 0]
            for link in self.links:
                pos = self._fk(z, link)
                for i in range(self.K):
                    o = ob[..., 4 * i: 4 * i + 4]
                    rows.append(torch.linalg.vector_norm(pos - o[..., :3], dim=-1) - o[..., 3] - rb)
        elif name == "LinearConstraints":  # inequalities/LinearConstraints.py:25-40, utils/utils.py:48-52
            rb = self._P(p, "r_body")[..., 0]
            for link in self.links:
                pos = self._fk(z, link)
                for i in range(self.K):
                    pl = self._P(p, f"lin_constrs_{i}")
                    d = torch.abs((pl[..., :3] * pos).sum(-1) + pl[..., 3]) / torch.linalg.vector_norm(pl[..., :3], dim=-1)
                    rows.append(d - rb)
        elif name == "SelfCollisionAvoidanceConstraints":  # :18-27
            rb = self._P(p, "r_body")[..., 0]
            for a, b in self.pairs:
                rows.append(torch.linalg.vector_norm(self._fk(z, a) - self._fk(z, b), dim=-1) - 2 * rb)
        elif name == "JointLimitConstraints":  # :20-31
            lo, hi = self._P(p, "lower_limits"), self._P(p, "upper_limits")
            for j in range(self.n):
                rows += [z[..., j] - lo[..., j], hi[..., j] - z[..., j]]
        elif name == "InputLimitConstraints":  # :18-29, u = z[-nu:]
            lo, hi = self._P(p, "lower_limits_u"), self._P(p, "upper_limits_u")
            for j in range(self.nu):
                rows += [z[..., self.nx + j] - lo[..., j], hi[..., j] - z[..., self.nx + j]]
        elif name == "VelLimitConstraints":  # :19-31, last two entries of qdot = z[n:nx]
            lo, hi = self._P(p, "lower_limits_vel"), self._P(p, "upper_limits_vel")
            for j in range(2):
                v = z[..., self.nx - 2 + j]
                rows += [v - lo[..., j], hi[..., j] - v]
        return rows

    def ineq(self, z, p) -> torch.Tensor:
        """h(z_k, p_k) >= 0, modules concatenated in config order (InequalityManager.py:25-33). [..., nh]"""
        rows = []
        for c in self.constraints:
            rows += self._module_rows(c, z, p)
        return torch.stack(rows, -1)

    LIMIT_MODULES = ("JointLimitConstraints", "InputLimitConstraints", "VelLimitConstraints")

    def ineq_general(self, z, p) -> torch.Tensor:
        """The rows of h that are not simple bounds on one variable (Radial, Linear, SelfCollision), in h order."""
        rows = []
        for c in self.constraints:
            if c not in self.LIMIT_MODULES:
                rows += self._module_rows(c, z, p)
        return torch.stack(rows, -1) if rows else torch.zeros(z.shape[:-1] + (0,), dtype=DT)

    def simple_bounds(self, P):
        """(lo, hi)[N, nvar]: the limit rows `z_j - lo_j >= 0`, `hi_j - z_j >= 0` of JointLimit / InputLimit / VelLimit
        (inequalities/{JointLimit,InputLimit,VelLimit}Constraints.py) intersected with the +-100 box of
        models/mpcModel.py:23-28,98-104 — the same feasible set, stated as bounds for solvers that treat bounds cheaply."""
        N = P.shape[0]
        lo = torch.full((N, self.nvar), -100.0, dtype=DT)
        hi = torch.full((N, self.nvar), 100.0, dtype=DT)
        for c in self.constraints:
            if c == "JointLimitConstraints":
                sl, names = slice(0, self.n), ("lower_limits", "upper_limits")
            elif c == "InputLimitConstraints":
                sl, names = slice(self.nx, self.nx + self.nu), ("lower_limits_u", "upper_limits_u")
            elif c == "VelLimitConstraints":
                sl, names = slice(self.nx - 2, self.nx), ("lower_limits_vel", "upper_limits_vel")
            else:
                continue
            lo[:, sl] = torch.maximum(lo[:, sl], self._P(P, names[0]))
            hi[:, sl] = torch.minimum(hi[:, sl], self._P(P, names[1]))
        return lo, hi

    def stage_cost(self, z, p) -> torch.Tensor:
        """J(z_k, p_k) (ObjectiveManager.py:28-46): modules + u' diag(wu) u."""
        J = torch.zeros(z.shape[:-1], dtype=DT)
        for o in self.objectives:
            if o == "GoalReaching":  # goal_reaching.py:19-34
                err = self._fk(z, self.rob["end_link"]) - self._P(p, "goal")
                J = J + (self._P(p, "wgoal") * err * err).sum(-1)
            elif o == "ConstraintAvoidance":  # constraint_avoidance.py:25-32: first row of every module, N times
                w = self._P(p, "wconstr")
                for i, c in enumerate(self.constraints):
                    rows = self._module_rows(c, z, p)
                    if rows:
                        # a zero weight switches the term off (all shipped configs but one module): evaluated as 0, not
                        # as 0/h, which is NaN when a bound-constrained solver puts the row exactly on h = 0
                        wi = w[..., i]
                        J = J + self.N * wi / torch.where(wi == 0, torch.ones_like(rows[0]), rows[0])
        u = z[..., self.nx:]
        return J + (self._P(p, "wu") * u * u).sum(-1)

    def f_cont(self, x, u):
        if self.rob["base_type"] == "holonomic":  # models/mpcModel.py:65-69
            return torch.cat([x[..., self.n:], u], -1)
        th, v, w = x[..., 2], x[..., 6], x[..., 7]  # models/diff_drive_mpc_model.py:24-41
        zz = torch.zeros_like(th)
        return torch.stack([torch.cos(th) * v, torch.sin(th) * v, w, zz, zz, zz, u[..., 0], u[..., 1]], -1)

    def step(self, x, u):
        """Phi: ERK2 (explicit midpoint [INFERRED]), Ts=dt, 5 nodes (models/mpcModel.py:118-120)."""
        h = self.dt / 5
        for _ in range(5):
            k1 = self.f_cont(x, u)
            x = x + h * self.f_cont(x + 0.5 * h * k1, u)
        return x

    # ---- whole-horizon functions on Z[N, nvar], P[N, npar] --------------------------------------
    def objective(self, Z, P):
        return self.stage_cost(Z, P).sum()

    def eq_defects(self, Z, xinit):
        """[x_1 - xinit ; x_{k+1} - Phi(x_k,u_k)]  (mpcModel.py:87-90,108)."""
        X, U = Z[:, : self.nx], Z[:, self.nx:]
        return torch.cat([(X[0] - xinit), (X[1:] - self.step(X[:-1], U[:-1])).reshape(-1)])

    def ineq_all(self, Z, P):
        return self.ineq(Z, P).reshape(-1)


def load_model(cfg_path: str, urdf_dir: str | None = None) -> OracleModel:
    import os
    with open(cfg_path) as f:
        setup = yaml.safe_load(f)
    up = setup["robot"]["urdf_file"]
    if not os.path.isabs(up):
        up = os.path.join(urdf_dir or os.path.join(os.path.dirname(cfg_path), "..", "assets"), up)
    with open(up) as f:
        return OracleModel(setup, f.read())
