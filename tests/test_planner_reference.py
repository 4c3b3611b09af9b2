"""Host side of the path against the REFERENCE'S OWN planner (boundary B1).

tests/golden/planner_reference.json holds what /root/reference/robotmpcs/planner/mpcPlanner.py itself hands to its solver
(`problem = {xinit, x0, all_parameters}`) and returns (`action`) when driven through its examples' call sequences with a
recording solver (tests/golden/make_planner_golden.py).  Here the same sequences go through
robot_mpcs_b200.planner.mpcPlanner.MPCPlanner; the buffers it hands to its solver must be identical, bit for bit.

CPU: the solver object is replaced by a recording fake (test-only; the product class has no such switch).
GPU: the real solver; numpy and device-resident planners, first call of every case (later calls depend on the plan).
"""
import importlib.util
import json
import os

import numpy as np
import pytest
import yaml

from conftest import CONFIG, GOLDEN, SOLVERS

_spec = importlib.util.spec_from_file_location("make_planner_golden", os.path.join(GOLDEN, "make_planner_golden.py"))
_gen = importlib.util.module_from_spec(_spec)
_spec.loader.exec_module(_gen)

CASES = json.load(open(os.path.join(GOLDEN, "planner_reference.json")))
NAMES = [k for k in CASES if isinstance(CASES[k], dict)]


def _args(args):
    out = []
    for a in args:
        if isinstance(a, list) and a and isinstance(a[0], list) and a[0] and a[0][0] == "sphere":
            out.append([_gen.Sphere(p, r) for _, p, r in a])
        else:
            out.append(np.array(a, dtype=float) if isinstance(a, list) else a)
    return out


class RecordingFake:
    """Stands where planner/solver.py:Solver stands; answers with the generator's synthetic plan."""
    def __init__(self, path):
        props = yaml.safe_load(open(os.path.join(path, "properties.yaml")))
        self.N = int(os.path.basename(path).split("_H")[1].split("_")[0])
        self.nx, self.nu = props["nx"], props["nu"]
        self.nvar = self.nx + self.nu + props["ns"]
        self.calls, self.control_mode = [], 0

    @classmethod
    def from_directory(cls, path, device=0, max_batch=1):
        return cls(path)

    def set_options(self, **kw):
        self.control_mode = kw.get("control_mode", self.control_mode)

    def solve_batch(self, xinit, x0, params):
        self.calls.append({"xinit": np.array(xinit).reshape(-1), "x0": np.array(x0).reshape(-1), "all_parameters": np.array(params).reshape(-1)})
        z = _gen.synthetic_plan(len(self.calls), self.N, self.nvar)
        # csrc/rmpc_solver.cu:write_outputs — "acc": u of stage 1; "vel": z_2[nvar-2nu : nvar-nu]
        act = z[0, self.nx:self.nx + self.nu] if self.control_mode == 0 else z[1, self.nvar - 2 * self.nu:self.nvar - self.nu]
        one = np.ones(1)
        return {"z": z[None], "action": act[None], "exitflag": np.ones(1, dtype=np.int32), "iters": one, "obj": one,
                "res_stat": one, "res_eq": one, "res_ineq": one, "res_comp": one}


def _replay(case, planner_cls, **kw):
    mpc = yaml.safe_load(open(os.path.join(CONFIG, case["config"])))["mpc"]
    pl = planner_cls(case["robot"], SOLVERS + os.sep, **kw, **mpc)
    pl.concretize()
    pl.reset()
    for meth, args in case["setup"]:
        getattr(pl, meth)(*_args(args))
    return pl


@pytest.mark.parametrize("name", NAMES)
def test_buffers_identical_to_reference_planner(name, monkeypatch, built):
    import robot_mpcs_b200.planner.mpcPlanner as mp
    monkeypatch.setattr(mp, "Solver", RecordingFake)
    case = CASES[name]
    pl = _replay(case, mp.MPCPlanner)
    actions = []
    for st in case["steps"]:
        if isinstance(st[0], str):
            getattr(pl, st[0])(*_args(st[1]))
            continue
        a, out, flag = pl.computeAction(*[np.array(x, dtype=float) for x in st])
        actions.append(a)
        assert flag == 1 and sorted(out.keys())[0] == "x" + "1".zfill(len(str(case["N"])))
    assert len(pl._solver.calls) == len(case["calls"])
    for mine, ref in zip(pl._solver.calls, case["calls"]):
        for k in ("xinit", "x0", "all_parameters"):
            assert np.array_equal(mine[k], np.array(ref[k])), (name, k)        # same numbers in the same places
    for a, r in zip(actions, case["actions"]):
        assert np.array_equal(np.asarray(a), np.array(r))


def test_reference_cannot_predict_dynamic_obstacles():
    """Recorded, not pinned: the reference's updateDynamicObstacles raises (no `dt` method, mpcPlanner.py:158), so the
    constant-acceleration prediction here follows its evident intent and is checked against the formula instead."""
    assert CASES["updateDynamicObstacles"].startswith("AttributeError")


@pytest.mark.gpu
@pytest.mark.parametrize("name", NAMES)
def test_real_solver_first_call_buffers(name, built):
    """The real planner (numpy B = 1, and device-resident B = 3 with the scenario broadcast): the buffers at the first
    solve equal the reference planner's; the solve succeeds or fails only with a solver status, never an exception."""
    import torch
    from robot_mpcs_b200.planner.mpcPlanner import MPCPlanner
    case = CASES[name]
    ref = case["calls"][0]
    first = next(st for st in case["steps"] if not isinstance(st[0], str))
    pl = _replay(case, MPCPlanner)
    a, out, flag = pl.computeAction(*[np.array(x, dtype=float) for x in first])
    assert np.array_equal(np.asarray(pl._params).reshape(-1), np.array(ref["all_parameters"]))
    assert np.array_equal(np.asarray(pl._x0).reshape(-1), np.array(ref["x0"]))
    assert np.array_equal(np.asarray(pl._xinit).reshape(-1), np.array(ref["xinit"]))
    assert np.asarray(a).shape == (pl._nu,) and isinstance(flag, int)
    B = 3
    pd = _replay(case, MPCPlanner, batch_size=B, device_resident=True)
    ad, zd, fd = pd.computeAction(*[torch.as_tensor(np.tile(np.array(x, dtype=float), (B, 1)), device="cuda:0") for x in first])
    P = pd._params3.cpu().numpy().reshape(B, -1)
    assert all(np.array_equal(P[b], np.array(ref["all_parameters"])) for b in range(B))
    assert all(np.array_equal(pd._x0[b].cpu().numpy().reshape(-1), np.array(ref["x0"])) for b in range(B))
    assert ad.shape == (B, pd._nu) and fd.shape == (B,)
    if flag == 1:
        assert np.allclose(ad.cpu().numpy(), np.asarray(a)[None], atol=1e-9)   # same problem, same kernel, any batch size
