"""ORACLE — test infrastructure only.  Solves the restated NLP (oracle/nlp.py) on the CPU in float64 with
scipy SLSQP (exact autograd gradients), optionally cross-checked by trust-constr, and certifies any
candidate solution with a solver-independent KKT check.  This is the stand-in named in BASELINE.md §2
for forcespro.nlp.Solver.solve (call site planner/mpcPlanner.py:262).  PARITY UNPINNED for the solve: FORCESPRO is proprietary and absent, the reference holds no solutions (the NLP it is given is pinned, see nlp.py).
"""
from __future__ import annotations

import numpy as np
import torch
from scipy.optimize import minimize, lsq_linear
from torch.func import jacrev, vmap

from .nlp import OracleModel, DT


def _t(a):
    return torch.as_tensor(np.asarray(a, dtype=np.float64), dtype=DT)


class Problem:
    """One NLP instance: xinit[nx], params[N*npar] flat stage-major (planner/mpcPlanner.py:91-104,246-250)."""

    def __init__(self, model: OracleModel, xinit, params):
        self.m = model
        self.xinit = _t(xinit)
        self.P = _t(params).reshape(model.N, model.npar)
        # x_1 is fixed by xinitidx: optimise over v = [u_1, z_2..z_N]
        self.nfree = model.N * model.nvar - model.nx

    def unpack(self, v: torch.Tensor) -> torch.Tensor:
        return torch.cat([self.xinit, v]).reshape(self.m.N, self.m.nvar)

    def pack(self, Z) -> np.ndarray:
        return np.asarray(Z, dtype=np.float64).reshape(-1)[self.m.nx:]

    def f(self, v):
        return self.m.objective(self.unpack(v), self.P)

    def c(self, v):
        return self.m.eq_defects(self.unpack(v), self.xinit)[self.m.nx:]

    def h(self, v):
        return self.m.ineq_all(self.unpack(v), self.P)

    def hg(self, v):
        """General (non-bound) rows only; the limit rows are handed to the solvers as bounds (nlp.py:simple_bounds)."""
        return self.m.ineq_general(self.unpack(v), self.P).reshape(-1)

    def bounds(self):
        lo, hi = self.m.simple_bounds(self.P)
        return lo.reshape(-1)[self.m.nx:].numpy(), hi.reshape(-1)[self.m.nx:].numpy()

    # numpy wrappers with autograd derivatives
    def f_np(self, v):
        return float(self.f(_t(v)))

    def g_np(self, v):
        t = _t(v).requires_grad_(True)
        (g,) = torch.autograd.grad(self.f(t), t)
        return g.numpy()

    def jac(self, fun, v):
        """Jacobian of pb.c or pb.h w.r.t. the free variables.  Both are stage-structured (c_k touches z_k and x_{k+1},
        h_k touches z_k), so the blocks come from per-stage autograd (vmap over the horizon) and are placed into the dense
        matrix by hand: the same numbers as autograd through the whole horizon, two orders of magnitude sooner."""
        if fun == self.c:
            return self.jac_c(v)
        if fun == self.h:
            return self.jac_h(v)
        if fun == self.hg:
            return self.jac_h(v, general=True)
        return torch.autograd.functional.jacobian(fun, _t(v), vectorize=True).numpy()

    def jac_c(self, v):
        m = self.m
        Z = self.unpack(_t(v))
        Jd = vmap(jacrev(lambda z: m.step(z[: m.nx], z[m.nx:])))(Z[:-1]).numpy()   # [N-1, nx, nvar]
        A = np.zeros(((m.N - 1) * m.nx, m.N * m.nvar))
        for k in range(m.N - 1):
            A[k * m.nx:(k + 1) * m.nx, k * m.nvar:(k + 1) * m.nvar] = -Jd[k]
            A[k * m.nx:(k + 1) * m.nx, (k + 1) * m.nvar:(k + 1) * m.nvar + m.nx] = np.eye(m.nx)
        return A[:, m.nx:]

    def jac_h(self, v, general=False):
        m = self.m
        Z = self.unpack(_t(v))
        fn = m.ineq_general if general else m.ineq
        Jh = vmap(jacrev(lambda z, p: fn(z, p)))(Z, self.P).numpy()                  # [N, nh, nvar]
        nh = Jh.shape[1]
        G = np.zeros((m.N * nh, m.N * m.nvar))
        for k in range(m.N):
            G[k * nh:(k + 1) * nh, k * m.nvar:(k + 1) * m.nvar] = Jh[k]
        return G[:, m.nx:]


def solve_slsqp(model: OracleModel, xinit, x0, params, maxiter=500, ftol=1e-14, rounds=3):
    """Returns dict(Z[N,nvar], obj, success, nit).  x0 is the [N*nvar] initial guess (mpcPlanner.py:228-236)."""
    pb = Problem(model, xinit, params)
    v = pb.pack(np.asarray(x0, dtype=np.float64).reshape(model.N, model.nvar))
    cons = [
        {"type": "eq", "fun": lambda v: pb.c(_t(v)).numpy(), "jac": lambda v: pb.jac(pb.c, v)},
        {"type": "ineq", "fun": lambda v: pb.hg(_t(v)).numpy(), "jac": lambda v: pb.jac(pb.hg, v)},
    ]
    # joint / input / velocity limit rows and the +-100 box of models/mpcModel.py:23-28,98-104 as bounds: the same
    # feasible set, and SLSQP's subproblem shrinks to the geometric rows
    lo, hi = pb.bounds()
    bounds = list(zip(lo.tolist(), hi.tolist()))
    nit, res = 0, None
    for _ in range(rounds):  # restarts polish the active-set solution
        res = minimize(pb.f_np, v, jac=pb.g_np, method="SLSQP", constraints=cons, bounds=bounds,
                       options={"maxiter": maxiter, "ftol": ftol})
        nit += res.nit
        if np.max(np.abs(res.x - v)) < 1e-12:
            v = res.x
            break
        v = res.x
    Z = pb.unpack(_t(v)).numpy()
    return {"Z": Z, "obj": float(res.fun), "success": bool(res.success), "nit": nit, "message": res.message}


def solve_trust_constr(model: OracleModel, xinit, x0, params, maxiter=3000):
    from scipy.optimize import NonlinearConstraint, Bounds
    pb = Problem(model, xinit, params)
    v = pb.pack(np.asarray(x0, dtype=np.float64).reshape(model.N, model.nvar))
    cons = [NonlinearConstraint(lambda v: pb.c(_t(v)).numpy(), 0, 0, jac=lambda v: pb.jac(pb.c, v)),
            NonlinearConstraint(lambda v: pb.hg(_t(v)).numpy(), 0, np.inf, jac=lambda v: pb.jac(pb.hg, v))]
    lo, hi = pb.bounds()
    res = minimize(pb.f_np, v, jac=pb.g_np, method="trust-constr", constraints=cons,
                   bounds=Bounds(lo, hi),
                   options={"maxiter": maxiter, "gtol": 1e-10, "xtol": 1e-14})
    return {"Z": pb.unpack(_t(res.x)).numpy(), "obj": float(res.fun), "success": bool(res.success), "nit": res.nit}


def kkt_certificate(model: OracleModel, xinit, params, Z, act_tol=1e-5):
    """Solver-independent check of a candidate Z[N,nvar]:
    eq   = max |dynamics / initial-state defect|
    ineq = max(0, -min h), bound violation folded in
    stat = min over (nu free, lam_A >= 0 on rows with h <= act_tol) of ||grad f + A' nu - G_A' lam||_inf,
           relative to max(1, ||grad f||_inf)
    """
    pb = Problem(model, xinit, params)
    Zt = np.asarray(Z, dtype=np.float64).reshape(model.N, model.nvar)
    eq0 = float(np.max(np.abs(Zt[0, : model.nx] - np.asarray(xinit, dtype=np.float64))))
    v = pb.pack(Zt)
    c = pb.c(_t(v)).numpy()
    h = pb.h(_t(v)).numpy()
    g = pb.g_np(v)
    A = pb.jac(pb.c, v)
    G = pb.jac(pb.h, v)
    act = np.where(h <= act_tol)[0]
    lb_act = np.where(v <= -100 + act_tol)[0]
    ub_act = np.where(v >= 100 - act_tol)[0]
    cols = [A.T, -G[act].T]
    nb = 0
    if len(lb_act):
        E = np.zeros((len(v), len(lb_act))); E[lb_act, np.arange(len(lb_act))] = -1; cols.append(E); nb += len(lb_act)
    if len(ub_act):
        E = np.zeros((len(v), len(ub_act))); E[ub_act, np.arange(len(ub_act))] = 1; cols.append(E); nb += len(ub_act)
    M = np.concatenate(cols, axis=1)
    lo = np.concatenate([-np.inf * np.ones(A.shape[0]), np.zeros(len(act) + nb)])
    sol = lsq_linear(M, -g, bounds=(lo, np.inf * np.ones_like(lo)), tol=1e-14, max_iter=500)
    r = M @ sol.x + g
    return {
        "eq": max(eq0, float(np.max(np.abs(c))) if c.size else 0.0),
        "ineq": max(0.0, float(-h.min())) + max(0.0, float(np.max(np.abs(v)) - 100.0)),
        "stat": float(np.max(np.abs(r)) / max(1.0, np.max(np.abs(g)))),
        "n_active": int(len(act) + nb),
        "obj": pb.f_np(v),
        "min_h": float(h.min()),
    }
