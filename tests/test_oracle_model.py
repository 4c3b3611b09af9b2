"""CPU tests: the oracle against the pins of SURVEY.md §3.5/§7, and the product's model functions (compiled for the
host through oracle/cpu_twin.cpp) against the oracle: values, and analytic derivatives vs oracle autograd."""
import json
import os

import numpy as np
import pytest
import torch
import yaml

from conftest import CONFIG, GOLDEN, SOLVERS
from oracle.nlp import load_model, DT
from oracle.build_twin import Twin
import ctypes

MODELS = {
    "pointRobotMpc.yaml": "pointRobot_n3_005_H20_noSlack",
    "boxerRadialMpc.yaml": "boxer_n3_01_H30_noSlack",
    "pandaMpc.yaml": "panda_n7_005_H20_noSlack",
    "boxerSupermarketMpc.yaml": "boxer_n3_01_H40_noSlack",
    "boxerMpc.yaml": "boxer_n3_01_H10_noSlack",
}


def test_fk_anchors():
    """SURVEY §3.5 anchors (computed there with an independent throwaway chain FK)."""
    m = load_model(os.path.join(CONFIG, "pandaMpc.yaml"))
    z = torch.zeros(7, dtype=DT)
    f = lambda q, l: m.fkm.fk(q, "panda_link0", l).numpy()
    assert np.allclose(f(z, "panda_link3"), [0, 0, 0.649], atol=1e-9)
    assert np.allclose(f(z, "panda_link5"), [0, 0, 1.033], atol=1e-9)
    assert np.allclose(f(z, "panda_link7"), [0.088, 0, 1.033], atol=1e-9)
    q = torch.tensor([0, 0, 0, -1.5708, 0, 1.8675, 0], dtype=DT)
    assert np.allclose(f(q, "panda_link5"), [0.4665, 0, 0.731499], atol=2e-6)
    assert np.allclose(f(q, "panda_link7"), [0.550655, 0, 0.757227], atol=2e-6)
    assert abs(np.linalg.norm(f(q, "panda_link3") - f(q, "panda_link7")) - 0.56119) < 2e-5
    mp = load_model(os.path.join(CONFIG, "pointRobotMpc.yaml"))
    assert np.allclose(mp.fkm.fk(torch.tensor([1.5, -2.0, 0.7], dtype=DT), "world", "base_link").numpy(), [1.5, -2.0, 0.05])
    mb = load_model(os.path.join(CONFIG, "boxerMpc.yaml"))
    p = mb.fkm.fk(torch.tensor([1.0, 2.0, 0.5], dtype=DT), "base_link", "ee_link").numpy()
    assert np.allclose(p, [1 + 0.4 * np.cos(0.5), 2 + 0.4 * np.sin(0.5), 0])


def test_param_maps_and_names():
    """paramMap tables and solver-directory names of SURVEY §3.1/§3.5."""
    m = load_model(os.path.join(CONFIG, "pointRobotMpc.yaml"))
    assert m.solver_name() == "pointRobot_n3_005_H20_noSlack" and m.npar == 30 and m.nh == 13
    assert m.paramMap["obst"] == [1, 2, 3, 4] and m.paramMap["wu"] == [17, 18, 19] and m.paramMap["wconstr"] == [26, 27, 28, 29]
    m = load_model(os.path.join(CONFIG, "pandaMpc.yaml"))
    assert m.solver_name() == "panda_n7_005_H20_noSlack" and m.npar == 50 and m.nh == 32
    assert m.paramMap["goal"] == [40, 41, 42] and m.paramMap["wgoal"] == [43, 44, 45]
    m = load_model(os.path.join(CONFIG, "boxerMpc.yaml"))
    assert m.solver_name() == "boxer_n3_01_H10_noSlack" and m.npar == 27 and (m.nx, m.nu) == (8, 2)
    assert m.paramMap["lin_constrs_0"] == [1, 2, 3, 4] and m.paramMap["wconstr"] == [23, 24, 25, 26]
    m = load_model(os.path.join(CONFIG, "boxerRadialMpc.yaml"))
    assert m.npar == 43 and m.nh == 15
    m = load_model(os.path.join(CONFIG, "boxerSupermarketMpc.yaml"))
    assert m.npar == 55 and m.nh == 18


@pytest.mark.skipif(not os.path.isdir("/root/reference"), reason="reference tree not present on this box")
def test_assets_match_reference_urdfs():
    """Our kinematic-skeleton URDFs give the same FK as the reference's full URDFs."""
    import yaml as _y
    from oracle.nlp import OracleModel
    for cfg, ref in (("pointRobotMpc.yaml", "pointRobot/pointRobot.urdf"), ("pandaMpc.yaml", "panda/panda.urdf"),
                     ("boxerMpc.yaml", "boxer/boxer_fk.urdf")):
        setup = _y.safe_load(open(os.path.join(CONFIG, cfg)))
        a = load_model(os.path.join(CONFIG, cfg))
        b = OracleModel(setup, open(os.path.join("/root/reference/examples/assets", ref)).read())
        g = torch.Generator().manual_seed(0)
        q = torch.rand(5, a.n, dtype=DT, generator=g) * 2 - 1
        for link in set(a.links + [a.rob["end_link"]]):
            assert torch.allclose(a.fkm.fk(q, a.rob["root_link"], link), b.fkm.fk(q, b.rob["root_link"], link), atol=1e-12)


def test_oracle_matches_survey_probe():
    """Config 1: objective 1361.93263, first action (1.0, -0.26100, 0) (SURVEY §7 step 0 / BASELINE.md §2)."""
    g = json.load(open(os.path.join(GOLDEN, "pointRobot_example.json")))
    assert abs(g["obj"] - 1361.93263) < 1e-4
    Z = np.array(g["Z"])
    assert np.allclose(Z[0, 6:9], [1.0, -0.26100, 0.0], atol=2e-5)
    assert np.allclose(Z[1, :6], [0.00625, -0.000326, 0, 0.15, -0.01305, 0], atol=2e-6)
    k = g["cases"][0]["kkt"]
    assert k["eq"] < 1e-9 and k["ineq"] < 1e-9 and k["stat"] < 1e-7
    assert g["cases"][0]["start"] == "x0"                                            # solved from the planner's own guess
    assert g["cases"][0]["trust_constr"]["max_abs_diff"] < 1e-4                      # a second solver from the same x0 agrees


def _twin(built, cfg):
    return Twin(os.path.join(built, MODELS[cfg]))


def _rand_stage(m, seed):
    g = np.random.default_rng(seed)
    z = g.uniform(-1, 1, m.nvar)
    p = g.uniform(0.2, 1.0, m.npar)
    for name, idx in m.paramMap.items():
        if name.startswith("lower"):
            p[idx] = -3 - g.uniform(0, 1, len(idx))
        if name.startswith("upper"):
            p[idx] = 3 + g.uniform(0, 1, len(idx))
        if name == "obst":
            p[idx] = np.tile([2.5, 2.0, 0.5, 0.3], len(idx) // 4) + g.uniform(-0.2, 0.2, len(idx))
        if name.startswith("lin_constrs"):
            p[idx] = [g.uniform(-1, 1), g.uniform(-1, 1), 0.0, 4.0]
        if name == "r_body":
            p[idx] = 0.2
    return z, p


@pytest.mark.parametrize("cfg", list(MODELS))
def test_model_values_match_oracle(built, cfg):
    m = load_model(os.path.join(CONFIG, cfg))
    tw = _twin(built, cfg)
    assert (tw.N, tw.nx, tw.nu, tw.nvar, tw.npar, tw.nh) == (m.N, m.nx, m.nu, m.nvar, m.npar, m.nh)
    assert yaml.safe_load(open(os.path.join(built, MODELS[cfg], "paramMap.yaml"))) == m.paramMap
    P = lambda a: a.ctypes.data_as(ctypes.c_void_p)
    for seed in range(5):
        z, p = _rand_stage(m, seed)
        f = ctypes.c_double()
        h = np.zeros(m.nh)
        tw.lib.twin_stage_values(P(z), P(p), ctypes.byref(f), P(h))
        zt, pt = torch.tensor(z, dtype=DT), torch.tensor(p, dtype=DT)
        assert np.allclose(h, m.ineq(zt, pt).numpy(), rtol=1e-10, atol=1e-10)  # URDF writes pi/2 to 1e-11; the generator snaps it
        assert abs(f.value - float(m.stage_cost(zt, pt))) <= 1e-8 * max(1, abs(f.value))
        xn = np.zeros(m.nx)
        AB = np.zeros((m.nx, m.nvar))
        tw.lib.twin_dynamics(P(z), P(xn), P(AB))
        assert np.allclose(xn, m.step(zt[: m.nx], zt[m.nx:]).numpy(), rtol=1e-12, atol=1e-12)
        J = torch.autograd.functional.jacobian(lambda v: m.step(v[: m.nx], v[m.nx:]), zt).numpy()
        assert np.allclose(AB, J, rtol=1e-10, atol=1e-12)


@pytest.mark.parametrize("cfg", list(MODELS))
def test_model_derivatives_match_autograd(built, cfg):
    """Analytic gradient / row Jacobian / exact Hessian of the fused stage function vs torch autograd on the oracle."""
    m = load_model(os.path.join(CONFIG, cfg))
    tw = _twin(built, cfg)
    mj = json.load(open(os.path.join(built, MODELS[cfg], "model.json")))
    geo_rows = [g["row"] for g in mj["geo"]]
    ngeo, nq = len(geo_rows), m.n
    P = lambda a: a.ctypes.data_as(ctypes.c_void_p)
    for seed in range(4):
        z, p = _rand_stage(m, 10 + seed)
        rng = np.random.default_rng(seed)
        rho = rng.uniform(0, 2, max(ngeo, 1))
        kappa = -rng.uniform(0, 2, max(ngeo, 1))
        f = ctypes.c_double()
        g = np.zeros(m.nvar); Hqq = np.zeros((nq, nq)); Hd = np.zeros(m.nvar); h = np.zeros(m.nh); Gg = np.zeros((max(ngeo, 1), nq))
        tw.lib.twin_stage_derivatives(P(z), P(p), P(rho), P(kappa), 1, ctypes.byref(f), P(g), P(Hqq), P(Hd), P(h), P(Gg))
        zt, pt = torch.tensor(z, dtype=DT), torch.tensor(p, dtype=DT)
        cost = lambda v: m.stage_cost(v, pt)
        g_ref = torch.autograd.functional.jacobian(cost, zt).numpy()
        assert np.abs(g - g_ref).max() <= 1e-9 * max(1.0, np.abs(g_ref).max())
        Jh = torch.autograd.functional.jacobian(lambda v: m.ineq(v, pt), zt).numpy()
        for i, r in enumerate(geo_rows):
            assert np.allclose(Gg[i], Jh[r, :nq], rtol=1e-9, atol=1e-11)
            assert np.allclose(Jh[r, nq:], 0)
        # exact Hessian of  cost + sum_r kappa_r h_r  (+ rho_r g g')  w.r.t. z
        kt = torch.tensor(kappa, dtype=DT)
        lag = lambda v: m.stage_cost(v, pt) + sum(kt[i] * m.ineq(v, pt)[r] for i, r in enumerate(geo_rows))
        H_ref = torch.autograd.functional.hessian(lag, zt).numpy()
        for i, r in enumerate(geo_rows):
            H_ref[:nq, :nq] += rho[i] * np.outer(Jh[r, :nq], Jh[r, :nq])
        H = np.diag(Hd)
        H[:nq, :nq] += Hqq
        assert np.abs(H - H_ref).max() <= 1e-8 * max(1.0, np.abs(H_ref).max()), np.abs(H - H_ref).max()
        # dynamics curvature
        nu = rng.uniform(-1, 1, m.nx)
        Q = np.zeros((m.nvar, m.nvar))
        tw.lib.twin_dyn_curvature(P(z), P(nu), P(Q))
        nut = torch.tensor(nu, dtype=DT)
        Q_ref = torch.autograd.functional.hessian(lambda v: (nut * m.step(v[: m.nx], v[m.nx:])).sum(), zt).numpy()
        assert np.allclose(Q, Q_ref, rtol=1e-9, atol=1e-11)


def test_fast_sincos():
    """The inline sincos of csrc/rmpc_model.cuh (same source on host and device) against libm over the fast-path range,
    at quadrant boundaries, and on the library fallback for huge arguments."""
    import ctypes
    from oracle.build_twin import Twin
    tw = Twin(os.path.join(SOLVERS, "boxer_n3_01_H30_noSlack"))
    rng = np.random.default_rng(0)
    x = np.concatenate([rng.uniform(-10, 10, 200000), rng.uniform(-1e5, 1e5, 200000), np.arange(-64, 65) * (np.pi / 4),
                        np.arange(-64, 65) * (np.pi / 4) + 1e-9, [0.0, 1e-300, -1e-8, 1e7, -3e9]])
    s, c = np.empty_like(x), np.empty_like(x)
    p = lambda a: a.ctypes.data_as(ctypes.c_void_p)
    tw.lib.twin_sincos(len(x), p(x), p(s), p(c))
    assert np.abs(s - np.sin(x)).max() < 4e-16 and np.abs(c - np.cos(x)).max() < 4e-16
