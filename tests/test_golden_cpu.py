"""CPU checks of the golden fixtures (tests/golden/*.json, every case solved by SLSQP from the planner's own x0):
  * the oracle reproduces its own stored certificates (the fixtures belong to the NLP of oracle/nlp.py as it is now);
  * the cross-check solver (trust-constr, same x0) is recorded for the first cases of every fixture;
  * the host twin of the solver algorithm (oracle/cpu_twin.cpp: the Ipm<> template the CUDA kernel compiles, G = 1)
    reaches the same basin-agreement floors the GPU test asks of the kernel — the algorithm is checked here without a GPU,
    the kernel itself in tests/test_gpu_parity.py::test_golden."""
import json
import os

import numpy as np
import pytest
import yaml

from conftest import CONFIG, GOLDEN, SOLVERS
from robot_mpcs_b200 import scenarios as sc

FLOORS = {"pointRobot_example": 1.0, "pointRobot_sweep": 0.97, "boxer_radial": 0.6, "boxer_supermarket": 0.9, "panda_example": 1.0,
          "panda_jitter": 0.97, "boxer_linear": 0.9, "boxer_vel": 0.93}
CFG_OF = {"pointRobot_n3_005_H20_noSlack": "pointRobotMpc.yaml", "boxer_n3_01_H30_noSlack": "boxerRadialMpc.yaml",
          "panda_n7_005_H20_noSlack": "pandaMpc.yaml", "boxer_n3_01_H40_noSlack": "boxerSupermarketMpc.yaml",
          "boxer_n3_01_H10_noSlack": "boxerMpc.yaml", "boxerVel_n3_01_H15_noSlack": "boxerVelMpc.yaml"}


def _load(name):
    g = json.load(open(os.path.join(GOLDEN, name + ".json")))
    pm = yaml.safe_load(open(os.path.join(SOLVERS, g["solver"], "paramMap.yaml")))
    return g, getattr(sc, g["generator"])(pm, **g["kwargs"])


def _valid(case):
    k = case["kkt"]
    return k["eq"] < 1e-6 and k["ineq"] < 1e-6 and k["stat"] < 1e-5


@pytest.mark.parametrize("name", list(FLOORS))
def test_fixture_is_from_x0_and_certified(built, name):
    from oracle.nlp import load_model
    from oracle.solve import kkt_certificate
    g, s = _load(name)
    assert "x0" in g["start"] and all(c["start"] == "x0" for c in g["cases"])
    assert [c["index"] for c in g["cases"]] == list(range(len(g["cases"])))          # contiguous, nothing skipped
    assert sum(_valid(c) for c in g["cases"]) >= 0.95 * len(g["cases"])
    model = load_model(os.path.join(CONFIG, CFG_OF[g["solver"]]))
    for c in g["cases"][:: max(1, len(g["cases"]) // 3)]:
        b = c["index"]
        k = kkt_certificate(model, s["xinit"][b], s["params"][b].reshape(-1), np.array(c["Z"]))
        assert abs(k["obj"] - c["obj"]) <= 1e-12 * abs(c["obj"]) and abs(k["eq"] - c["kkt"]["eq"]) < 1e-9 and k["n_active"] == c["kkt"]["n_active"]
    tc = [c["trust_constr"] for c in g["cases"] if "trust_constr" in c]
    if tc:                                                                          # a second solver from the same x0
        assert all(t["kkt"]["eq"] < 1e-6 and t["kkt"]["ineq"] < 1e-6 for t in tc)


@pytest.mark.parametrize("name", list(FLOORS))
def test_twin_reaches_basin_agreement_floor(built, name):
    from oracle.build_twin import Twin
    g, s = _load(name)
    r = Twin(os.path.join(SOLVERS, g["solver"])).solve(s["xinit"], s["x0"], s["params"])
    same = valid = 0
    for c in g["cases"]:
        if not _valid(c):
            continue
        valid += 1
        b = c["index"]
        zo = np.array(c["Z"])
        assert r["exitflag"][b] == 1, (name, b, r["exitflag"][b])
        rel = np.abs(r["z"][b] - zo).max() / np.abs(zo).max()
        same += bool(rel < 1e-4 and abs(r["obj"][b] - c["obj"]) <= 1e-5 * abs(c["obj"]))
    assert same >= FLOORS[name] * valid, (name, same, valid)


def test_twin_refuses_non_finite_inputs_before_the_first_iteration(built):
    """The algorithm's host build (same template as the kernel): NaN / Inf in xinit, x0 or the parameters -> exit flag -6, no iteration,
    neighbours unaffected (the GPU form of this test is tests/test_gpu_features.py::test_nan_and_inf_inputs_...)."""
    import yaml
    from oracle.build_twin import Twin
    from robot_mpcs_b200 import scenarios as sc
    for sdir, gen in [("boxer_n3_01_H30_noSlack", sc.boxer_radial), ("pointRobot_n3_005_H20_noSlack", sc.point_robot_sweep),
                      ("panda_n7_005_H20_noSlack", sc.panda)]:
        pm = yaml.safe_load(open(os.path.join(SOLVERS, sdir, "paramMap.yaml")))
        s = gen(pm, 12)
        tw = Twin(os.path.join(SOLVERS, sdir))
        clean = tw.solve(s["xinit"], s["x0"], s["params"])
        x, x0, P = s["xinit"].copy(), s["x0"].copy(), s["params"].copy()
        x[1, 0] = np.nan
        x0[2, 3, 1] = np.nan
        P[3, :, pm["goal"][0]] = np.inf
        P[4, 5, pm["wu"][0]] = np.nan
        x[5, 1] = -np.inf
        r = tw.solve(x, x0, P)
        assert (r["exitflag"][1:6] == -6).all() and (r["iters"][1:6] == 0).all(), (sdir, r["exitflag"], r["iters"])
        keep = np.r_[0, 6:12]
        assert np.array_equal(r["exitflag"][keep], clean["exitflag"][keep]) and np.array_equal(r["z"][keep], clean["z"][keep])
