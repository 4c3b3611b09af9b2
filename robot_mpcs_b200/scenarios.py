"""Synthetic scenario batches for the BASELINE.json configs.  Inputs only: every array is
reference-shaped (xinit[B,nx], x0[B,N,nvar], params[B,N,npar], float64) and filled through the paramMap, the
way planner/mpcPlanner.py:83-236 of the reference fills its buffers.  Seeds are fixed so tests, the
oracle, the CPU twin and the GPU see identical numbers.
"""
from __future__ import annotations

import math
from typing import Dict, List

import numpy as np


def _fill(params: np.ndarray, pmap: Dict[str, List[int]], name: str, val):
    """params[B,N,npar]; val broadcastable to [B,N,len] or [B,len]."""
    idx = pmap[name]
    v = np.asarray(val, dtype=np.float64)
    if v.ndim == 1:
        v = v[None, None, :]
    elif v.ndim == 2:
        v = v[:, None, :]
    params[:, :, idx] = v


def _rng(seed):
    return np.random.Generator(np.random.PCG64(seed))


def point_robot_example(pmap, N=20, npar=30):
    """Config 1, exact constants of examples/pointRobot_example.py:31-65 + config/pointRobotMpc.yaml."""
    B = 1
    P = np.zeros((B, N, npar))
    _fill(P, pmap, "r_body", [0.3])
    _fill(P, pmap, "obst", [4.0, -0.5, 0.0, 1.0])
    _fill(P, pmap, "lower_limits", [-10.0] * 3)
    _fill(P, pmap, "upper_limits", [10.0] * 3)
    _fill(P, pmap, "lower_limits_u", [-1.0, -1.0, -15.0])
    _fill(P, pmap, "upper_limits_u", [1.0, 1.0, 15.0])
    _fill(P, pmap, "wu", [0.1] * 3)
    _fill(P, pmap, "goal", [8.2, -0.2, 0.0])
    _fill(P, pmap, "wgoal", [1.0] * 3)
    _fill(P, pmap, "wconstr", [0.5, 0, 0, 0])
    xinit = np.array([[0, 0, 0, 0.1, 0, 0]], dtype=np.float64)
    x0 = np.zeros((B, N, 9))
    x0[:, :, :6] = xinit[:, None, :]
    return {"xinit": xinit, "x0": x0, "params": P}


def point_robot_sweep(pmap, B, N=20, npar=30, seed=3):
    """Config 5."""
    g = _rng(seed)
    start = g.uniform(-2, 2, (B, 2))
    vel = g.uniform(-0.2, 0.2, (B, 2))
    goal = g.uniform(-8, 8, (B, 2))
    d = goal - start
    L = np.linalg.norm(d, axis=1, keepdims=True) + 1e-9
    perp = np.stack([-d[:, 1], d[:, 0]], 1) / L
    t = g.uniform(0.3, 0.7, (B, 1))
    off = g.uniform(0.3, 1.5, (B, 1)) * np.where(g.uniform(size=(B, 1)) < 0.5, -1.0, 1.0)
    r = g.uniform(0.5, 1.0, (B, 1))
    c = start + t * d + perp * off
    # start clearance >= 0.2 : push the obstacle away along the start->centre ray if needed
    r_body = 0.3
    v = c - start
    dist = np.linalg.norm(v, axis=1, keepdims=True)
    need = r + r_body + 0.2 + 0.05  # z offset 0.05 only helps
    c = np.where(dist < need, start + v / np.maximum(dist, 1e-9) * need, c)
    P = np.zeros((B, N, npar))
    _fill(P, pmap, "r_body", [r_body])
    _fill(P, pmap, "obst", np.concatenate([c, np.zeros((B, 1)), r], 1))
    _fill(P, pmap, "lower_limits", [-10.0] * 3)
    _fill(P, pmap, "upper_limits", [10.0] * 3)
    _fill(P, pmap, "lower_limits_u", [-1.0, -1.0, -15.0])
    _fill(P, pmap, "upper_limits_u", [1.0, 1.0, 15.0])
    _fill(P, pmap, "wu", [0.1] * 3)
    _fill(P, pmap, "goal", np.concatenate([goal, np.zeros((B, 1))], 1))
    _fill(P, pmap, "wgoal", [1.0] * 3)
    _fill(P, pmap, "wconstr", [0.5, 0, 0, 0])
    xinit = np.concatenate([start, np.zeros((B, 1)), vel, np.zeros((B, 1))], 1)
    x0 = np.zeros((B, N, 9))
    x0[:, :, :6] = xinit[:, None, :]
    return {"xinit": xinit, "x0": x0, "params": P}


def _boxer_start(g, B):
    xy = g.uniform(-1, 1, (B, 2))
    th = g.uniform(-math.pi, math.pi, (B, 1))
    v = g.uniform(0, 0.5, (B, 1))
    return np.concatenate([xy, th, np.zeros((B, 3)), v, np.zeros((B, 1))], 1)


def _boxer_common(P, pmap, B):
    _fill(P, pmap, "r_body", [0.6])
    _fill(P, pmap, "lower_limits", [-10.0] * 3)
    _fill(P, pmap, "upper_limits", [10.0] * 3)
    _fill(P, pmap, "lower_limits_u", [-10.0] * 2)
    _fill(P, pmap, "upper_limits_u", [10.0] * 2)
    _fill(P, pmap, "wu", [0.01] * 2)
    _fill(P, pmap, "wgoal", [1.0] * 3)
    _fill(P, pmap, "wconstr", [0.0] * 4)


def boxer_radial(pmap, B, N=30, npar=43, K=5, seed=0):
    """Config 2: boxer N=30, 5 sphere obstacles (constants examples/boxer_example.py:53,81-91)."""
    g = _rng(seed)
    xinit = _boxer_start(g, B)
    ee = xinit[:, :2] + 0.4 * np.stack([np.cos(xinit[:, 2]), np.sin(xinit[:, 2])], 1)
    goal = np.array([7.2, -2.2]) + g.uniform(-1, 1, (B, 2))
    obst = np.zeros((B, K, 4))
    for k in range(K):
        c = g.uniform([1, -3], [6, 1], (B, 2))
        r = g.uniform(0.2, 0.8, (B,))
        for _ in range(50):
            bad = np.linalg.norm(c - ee, axis=1) - r - 0.6 < 0.2
            if not bad.any():
                break
            c[bad] = g.uniform([1, -3], [6, 1], (int(bad.sum()), 2))
        bad = np.linalg.norm(c - ee, axis=1) - r - 0.6 < 0.2
        c[bad] = np.array([50.0, 50.0])
        obst[:, k, :2] = c
        obst[:, k, 3] = r
    P = np.zeros((B, N, npar))
    _boxer_common(P, pmap, B)
    _fill(P, pmap, "obst", obst.reshape(B, 4 * K))
    _fill(P, pmap, "goal", np.concatenate([goal, np.zeros((B, 1))], 1))
    x0 = np.zeros((B, N, 10))
    x0[:, :, :8] = xinit[:, None, :]
    return {"xinit": xinit, "x0": x0, "params": P}


def boxer_supermarket(pmap, B, N=40, npar=55, K=8, seed=2):
    """Config 4: boxer with K=8 stage-varying half-planes (stands in for per-stage free-space decomposition,
    examples/boxer_example_supermarket.py:246-256)."""
    g = _rng(seed)
    xinit = _boxer_start(g, B)
    ee = xinit[:, :2] + 0.4 * np.stack([np.cos(xinit[:, 2]), np.sin(xinit[:, 2])], 1)
    ang = g.uniform(0, 2 * math.pi, (B, 1))
    goal = xinit[:, :2] + g.uniform(1, 3, (B, 1)) * np.concatenate([np.cos(ang), np.sin(ang)], 1)
    P = np.zeros((B, N, npar))
    _boxer_common(P, pmap, B)
    _fill(P, pmap, "goal", np.concatenate([goal, np.zeros((B, 1))], 1))
    for i in range(K):
        a = g.uniform(0, 2 * math.pi, (B, 1))
        nrm = np.concatenate([np.cos(a), np.sin(a)], 1)  # unit normal
        clear = g.uniform(0.8, 3.0, (B, 1))
        d0 = clear - (nrm * ee).sum(1, keepdims=True)  # a.ee + d = clear > 0
        d = d0[:, None, :] + g.normal(0, 0.02, (B, N, 1))
        plane = np.concatenate([np.broadcast_to(nrm[:, None, :], (B, N, 2)), np.zeros((B, N, 1)), d], 2)
        _fill(P, pmap, f"lin_constrs_{i}", plane)
    x0 = np.zeros((B, N, 10))
    x0[:, :, :8] = xinit[:, None, :]
    return {"xinit": xinit, "x0": x0, "params": P}


def boxer_linear(pmap, B, N=10, npar=27, seed=5):
    """The reference's shipped boxer config (config/boxerMpc.yaml: N=10, one half-plane from LinearConstraints, K=1): the
    plane is the same for every stage of a scenario here (examples/boxer_example.py:93 uses one filler plane); the start
    ee point sits U(0.8, 2.0) on its positive side and the goal U(1, 3) away in a random direction, so the plane is
    active for roughly a third of the scenarios."""
    g = _rng(seed)
    xinit = _boxer_start(g, B)
    ee = xinit[:, :2] + 0.4 * np.stack([np.cos(xinit[:, 2]), np.sin(xinit[:, 2])], 1)
    ang = g.uniform(0, 2 * math.pi, (B, 1))
    goal = xinit[:, :2] + g.uniform(1, 3, (B, 1)) * np.concatenate([np.cos(ang), np.sin(ang)], 1)
    P = np.zeros((B, N, npar))
    _boxer_common(P, pmap, B)
    _fill(P, pmap, "goal", np.concatenate([goal, np.zeros((B, 1))], 1))
    a = g.uniform(0, 2 * math.pi, (B, 1))
    nrm = np.concatenate([np.cos(a), np.sin(a)], 1)
    clear = g.uniform(0.8, 2.0, (B, 1))
    d0 = clear - (nrm * ee).sum(1, keepdims=True)
    _fill(P, pmap, "lin_constrs_0", np.concatenate([nrm, np.zeros((B, 1)), d0], 1))
    x0 = np.zeros((B, N, 10))
    x0[:, :, :8] = xinit[:, None, :]
    return {"xinit": xinit, "x0": x0, "params": P}


def boxer_vel(pmap, B, N=15, npar=36, K=2, seed=7):
    """config/boxerVelMpc.yaml: boxer with VelLimitConstraints on (v, omega) (inequalities/VelLimitConstraints.py:19-31) and
    control_mode vel (planner/mpcPlanner.py:275-278).  Goal 3-5 m away, forward speed limited to 1 m/s and turn rate to
    1 rad/s, so both velocity limits become active; two sphere obstacles as in boxer_radial."""
    g = _rng(seed)
    xinit = _boxer_start(g, B)
    ee = xinit[:, :2] + 0.4 * np.stack([np.cos(xinit[:, 2]), np.sin(xinit[:, 2])], 1)
    ang = g.uniform(0, 2 * math.pi, (B, 1))
    goal = xinit[:, :2] + g.uniform(3, 5, (B, 1)) * np.concatenate([np.cos(ang), np.sin(ang)], 1)
    obst = np.zeros((B, K, 4))
    for k in range(K):
        c = ee + g.uniform(-4, 4, (B, 2))
        r = g.uniform(0.2, 0.6, (B,))
        for _ in range(50):
            bad = np.linalg.norm(c - ee, axis=1) - r - 0.6 < 0.2
            if not bad.any():
                break
            c[bad] = ee[bad] + g.uniform(-4, 4, (int(bad.sum()), 2))
        bad = np.linalg.norm(c - ee, axis=1) - r - 0.6 < 0.2
        c[bad] = np.array([50.0, 50.0])
        obst[:, k, :2] = c
        obst[:, k, 3] = r
    P = np.zeros((B, N, npar))
    _fill(P, pmap, "r_body", [0.6])
    _fill(P, pmap, "obst", obst.reshape(B, 4 * K))
    _fill(P, pmap, "lower_limits", [-10.0] * 3)
    _fill(P, pmap, "upper_limits", [10.0] * 3)
    _fill(P, pmap, "lower_limits_u", [-10.0] * 2)
    _fill(P, pmap, "upper_limits_u", [10.0] * 2)
    _fill(P, pmap, "lower_limits_vel", [-0.3, -1.0])
    _fill(P, pmap, "upper_limits_vel", [1.0, 1.0])
    _fill(P, pmap, "wu", [0.01] * 2)
    _fill(P, pmap, "goal", np.concatenate([goal, np.zeros((B, 1))], 1))
    _fill(P, pmap, "wgoal", [1.0] * 3)
    _fill(P, pmap, "wconstr", [0.0] * 5)
    x0 = np.zeros((B, N, 10))
    x0[:, :, :8] = xinit[:, None, :]
    return {"xinit": xinit, "x0": x0, "params": P}


PANDA_LIMITS = np.array([[-2.8973, 2.8973], [-1.7628, 1.7628], [-2.8973, 2.8973], [-3.0718, -0.0698],
                         [-2.8973, 2.8973], [-0.0175, 3.7525], [-2.8973, 2.8973]])
PANDA_LIMITS_U = np.array([[-1, 1], [-1, 1], [-15, 15], [-15, 15], [-7.5, 7.5], [-10, 10], [-12.5, 12.5]], dtype=np.float64)


def panda(pmap, B, N=20, npar=50, seed=1, jitter=True):
    """Config 3: pandaMpc.yaml + constants examples/panda_example.py:37,53-78.  jitter=False gives the
    example's own single scenario."""
    g = _rng(seed)
    q0 = np.tile(np.median(PANDA_LIMITS, axis=1), (B, 1))
    goal = np.tile(np.array([0.1, -0.6, 0.4]), (B, 1))
    oc = np.tile(np.array([0.5, -0.3, 0.3]), (B, 1))
    if jitter:
        q0 = q0 + g.uniform(-0.3, 0.3, (B, 7))
        q0 = np.clip(q0, PANDA_LIMITS[:, 0] + 0.05, PANDA_LIMITS[:, 1] - 0.05)
        goal = goal + g.uniform(-0.15, 0.15, (B, 3))
        oc = oc + g.uniform(-0.1, 0.1, (B, 3))
    if jitter:
        # "reject if any h < 0.02 at start" — a measured state that already touches the obstacle makes
        # the NLP infeasible (x_1 = xinit is fixed and the rows apply at stage 1).  Scenarios whose draw is fine keep it;
        # the others get a new obstacle centre from a second, equally seeded stream.
        import os
        from robot_mpcs_b200.utils.urdf import KinematicTree
        links = ["panda_link3", "panda_link5", "panda_link7"]
        with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "assets", "panda.urdf")) as f:
            pos = KinematicTree(f.read()).positions(q0, "panda_link0", "panda_link7", links)
        g2 = _rng(seed + 1000)
        for _ in range(50):
            clear = np.min([np.linalg.norm(pos[l] - oc, axis=1) for l in links], axis=0) - 0.1 - 0.14
            bad = clear < 0.02
            if not bad.any():
                break
            oc[bad] = np.array([0.5, -0.3, 0.3]) + g2.uniform(-0.1, 0.1, (int(bad.sum()), 3))
    P = np.zeros((B, N, npar))
    _fill(P, pmap, "r_body", [0.14])
    _fill(P, pmap, "obst", np.concatenate([oc, 0.1 * np.ones((B, 1))], 1))
    _fill(P, pmap, "lower_limits", PANDA_LIMITS[:, 0])
    _fill(P, pmap, "upper_limits", PANDA_LIMITS[:, 1])
    _fill(P, pmap, "lower_limits_u", PANDA_LIMITS_U[:, 0])
    _fill(P, pmap, "upper_limits_u", PANDA_LIMITS_U[:, 1])
    _fill(P, pmap, "wu", [0.1] * 7)
    _fill(P, pmap, "goal", goal)
    _fill(P, pmap, "wgoal", [3.0] * 3)
    _fill(P, pmap, "wconstr", [0.5, 0, 0, 0])
    xinit = np.concatenate([q0, np.zeros((B, 7))], 1)
    x0 = np.zeros((B, N, 21))
    x0[:, :, :14] = xinit[:, None, :]
    return {"xinit": xinit, "x0": x0, "params": P}


def lidar_scene(B, N, P=64, seed=0):
    """Synthetic input of the free-space decomposition (stands in for urdfenvs' Lidar sensor, boxer_example.py:153-161,
    which is unavailable here): per problem a robot pose, N reference states along a gently curving path from it, and
    P lidar returns at `range_i` along P equally spaced rays from the lidar position (range ~ U(0.3, 6), so some fall
    outside max_radius = 5), at the lidar height 0.02.  Returns points[B,P,3] (absolute) and ref[B,N,3] = (x, y, theta)."""
    g = np.random.default_rng(seed)
    x0 = g.uniform(-1, 1, (B, 2))
    th0 = g.uniform(-math.pi, math.pi, B)
    v = g.uniform(0.0, 0.5, B)
    om = g.uniform(-0.5, 0.5, B)
    t = np.arange(N)[None, :] * 0.1
    th = th0[:, None] + om[:, None] * t
    ref = np.stack([x0[:, 0:1] + v[:, None] * t * np.cos(th), x0[:, 1:2] + v[:, None] * t * np.sin(th), th], axis=2)
    ang = th0[:, None] + np.arange(P)[None, :] * (2 * math.pi / P)
    rng_ = g.uniform(0.3, 6.0, (B, P))
    lp = x0 + 0.4 * np.stack([np.cos(th0), np.sin(th0)], axis=1)
    pts = np.stack([lp[:, 0:1] + rng_ * np.cos(ang), lp[:, 1:2] + rng_ * np.sin(ang), np.full((B, P), 0.02)], axis=2)
    return np.ascontiguousarray(pts), np.ascontiguousarray(ref)
