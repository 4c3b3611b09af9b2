"""Golden vectors for the host side of the path (boundary B1), produced by the REFERENCE'S OWN
planner class: /root/reference/robotmpcs/planner/mpcPlanner.py is imported here and driven through the call sequence of
its examples; what it hands to its solver (`problem = {xinit, x0, all_parameters}`, mpcPlanner.py:246-250) and what it
returns (`action`, mpcPlanner.py:275-281) is recorded.

The planner's third-party imports are absent offline and are not needed for this: `forcespro`, `casadi` and
`forwardkinematics` are stubbed with empty modules; `forcespro.nlp.Solver.from_directory` returns a recording solver
whose `solve(problem)` answers with a deterministic synthetic plan (so that the next call exercises `shiftHorizon`).
The solver directories (paramMap.yaml / properties.yaml, the on-disk contract of mpcModel.py:132-136) are this repo's.

Run in the build container only (needs /root/reference):  python tests/golden/make_planner_golden.py
-> tests/golden/planner_reference.json
"""
import json
import os
import sys
import types

import numpy as np
import yaml

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
SOLVERS = os.path.join(ROOT, "robot_mpcs_b200", "solvers") + os.sep
CONFIG = os.path.join(ROOT, "robot_mpcs_b200", "config")


def synthetic_plan(call, N, nvar):
    """The recording solver's answer: a smooth, call-dependent plan (any numbers do; the test feeds the same ones)."""
    k = np.arange(N)[:, None]
    j = np.arange(nvar)[None, :]
    return np.sin(0.37 * (k + 1) + 0.11 * j + 0.5 * call) + 0.01 * k * j


class RecordingSolver:
    def __init__(self, N, nvar):
        self.N, self.nvar, self.calls = N, nvar, []

    def solve(self, problem):
        self.calls.append({k: np.array(problem[k], dtype=float).reshape(-1).tolist() for k in ("xinit", "x0", "all_parameters")})
        z = synthetic_plan(len(self.calls), self.N, self.nvar)
        w = len(str(self.N))
        return {"x" + str(k + 1).zfill(w): z[k] for k in range(self.N)}, 1, None


def install_stubs():
    fp = types.ModuleType("forcespro")
    fp.nlp = types.ModuleType("forcespro.nlp")

    class Solver:
        @staticmethod
        def from_directory(path):
            props = yaml.safe_load(open(os.path.join(path, "properties.yaml")))
            N = int(os.path.basename(path).split("_H")[1].split("_")[0])
            return RecordingSolver(N, props["nx"] + props["nu"] + props["ns"])
    fp.nlp.Solver = Solver
    sys.modules["forcespro"] = fp
    sys.modules["forcespro.nlp"] = fp.nlp
    sys.modules["casadi"] = types.ModuleType("casadi")
    for name, attr in (("forwardkinematics", None), ("forwardkinematics.fksCommon", None), ("forwardkinematics.fksCommon.fk", "ForwardKinematics"),
                       ("forwardkinematics.urdfFks", None), ("forwardkinematics.urdfFks.generic_urdf_fk", "GenericURDFFk")):
        m = types.ModuleType(name)
        if attr:
            setattr(m, attr, type(attr, (), {}))
        sys.modules[name] = m


class Sphere:  # what the examples pass: mpscenes SphereObstacle, used through position()/radius()/dimension()
    def __init__(self, pos, r):
        self._p, self._r = list(pos), r

    def position(self):
        return self._p

    def radius(self):
        return self._r

    def dimension(self):
        return len(self._p)


def cases():
    """name -> (robotType, config yaml, setup calls [(method, args)], computeAction argument tuples).  Values follow
    examples/pointRobot_example.py:31-64, boxer_example.py:46-93, panda_example.py:30-83, boxer_example_supermarket.py."""
    rng = np.random.default_rng(5)
    planes = rng.normal(size=(3, 40, 8, 4)).round(6)
    return {
        "pointRobot": ("pointRobot", "pointRobotMpc.yaml", [
            ("setRadialConstraints", ([("sphere", [4.0, -0.5, 0.0], 1.0)], 0.3)),
            ("setSelfCollisionAvoidanceConstraints", (0.3,)),
            ("setJointLimits", ([[-10, -10, -10], [10, 10, 10]],)),
            ("setInputLimits", ([[-1, -1, -15], [1, 1, 15]],)),
            ("setGoalReaching", ([8.2, -0.2],)),
            ("setConstraintAvoidance", ()),
        ], [([0.0, 0.0, 0.0], [0.1, 0.0, 0.0]),
            ([0.05, -0.01, 0.0], [0.3, -0.1, 0.0]),
            ([0.1, -0.02, 0.0], [0.5, -0.2, 0.0])]),
        "boxer_radial": ("boxer", "boxerRadialMpc.yaml", [
            ("setRadialConstraints", ([("sphere", [3.0, -1.0, 0.0], 0.5), ("sphere", [5.0, 0.5, 0.0], 0.3), ("sphere", [1.5, 0.8, 0.0], 0.25),
                                      ("sphere", [4.2, -2.7, 0.0], 0.7), ("sphere", [2.2, -2.0, 0.0], 0.4)], 0.6)),   # exactly number_obstacles: the
            # reference's padding object lacks dimension() (mpcPlanner.py:18-26 vs :129), fewer obstacles raise there
            ("setSelfCollisionAvoidanceConstraints", (0.6,)),
            ("setJointLimits", ([[-10, -10, -10], [10, 10, 10]],)),
            ("setInputLimits", ([[-10, -10], [10, 10]],)),
            ("setGoalReaching", ([7.2, -2.2],)),
            ("setConstraintAvoidance", ()),
        ], [([0.0, 0.0, 0.3], [0.0, 0.0, 0.0], [0.2, 0.0]),
            ([0.02, 0.006, 0.31], [0.2, 0.06, 0.1], [0.25, 0.1])]),
        "panda": ("panda", "pandaMpc.yaml", [
            ("setRadialConstraints", ([("sphere", [0.5, -0.3, 0.3], 0.1)], 0.14)),
            ("setSelfCollisionAvoidanceConstraints", (0.14,)),
            ("setJointLimits", ([[-2.8973, -1.7628, -2.8973, -3.0718, -2.8973, -0.0175, -2.8973],
                                 [2.8973, 1.7628, 2.8973, -0.0698, 2.8973, 3.7525, 2.8973]],)),
            ("setInputLimits", ([[-15, -7.5, -10, -12.5, -15, -20, -20], [15, 7.5, 10, 12.5, 15, 20, 20]],)),
            ("setGoalReaching", ([0.1, -0.6, 0.4],)),
            ("setConstraintAvoidance", ()),
        ], [([0.0, -0.5, 0.0, -1.8, 0.0, 1.5, 0.2], [0.0] * 7),
            ([0.01, -0.5, 0.0, -1.79, 0.0, 1.5, 0.21], [0.1, 0.0, 0.0, 0.1, 0.0, 0.0, 0.1])]),
        # LinearConstraints are rewritten before every solve (boxer_example.py:203); previous_plan warm start
        "boxer_supermarket": ("boxer", "boxerSupermarketMpc.yaml", [
            ("setLinearConstraints", (planes[0].tolist(), 0.6)),
            ("setSelfCollisionAvoidanceConstraints", (0.6,)),
            ("setJointLimits", ([[-10, -10, -10], [10, 10, 10]],)),
            ("setInputLimits", ([[-10, -10], [10, 10]],)),
            ("setGoalReaching", ([2.0, 1.0],)),
            ("setConstraintAvoidance", ()),
        ], [([0.0, 0.0, 0.3], [0.0, 0.0, 0.0], [0.2, 0.0]),
            ("setLinearConstraints", (planes[1].tolist(), 0.6)),
            ([0.02, 0.006, 0.31], [0.2, 0.06, 0.1], [0.25, 0.1]),
            ("setLinearConstraints", (planes[2].tolist(), 0.6)),
            ([0.04, 0.012, 0.32], [0.2, 0.06, 0.1], [0.3, 0.1])]),
    }


def materialise(args):
    out = []
    for a in args:
        if isinstance(a, list) and a and isinstance(a[0], tuple) and a[0][0] == "sphere":
            out.append([Sphere(p, r) for _, p, r in a])
        else:
            out.append(np.array(a, dtype=float) if isinstance(a, list) else a)
    return out


def main():
    install_stubs()
    # the package's __init__ asks importlib.metadata for an installed distribution: load the two source files directly
    import importlib.util
    for pkg in ("robotmpcs", "robotmpcs.models", "robotmpcs.planner"):
        sys.modules[pkg] = types.ModuleType(pkg)
    for modname, path in (("robotmpcs.models.mpcBase", "/root/reference/robotmpcs/models/mpcBase.py"),
                          ("robotmpcs.planner.mpcPlanner", "/root/reference/robotmpcs/planner/mpcPlanner.py")):
        spec = importlib.util.spec_from_file_location(modname, path)
        mod = importlib.util.module_from_spec(spec)
        sys.modules[modname] = mod
        spec.loader.exec_module(mod)
    MPCPlanner = sys.modules["robotmpcs.planner.mpcPlanner"].MPCPlanner  # the reference's own class
    golden = {}
    for name, (robot, cfg, setup, steps) in cases().items():
        mpc = yaml.safe_load(open(os.path.join(CONFIG, cfg)))["mpc"]
        pl = MPCPlanner(robot, SOLVERS, **mpc)
        pl.concretize()
        pl.reset()
        for meth, args in setup:
            getattr(pl, meth)(*materialise(args))
        actions = []
        for st in steps:
            if isinstance(st[0], str):
                getattr(pl, st[0])(*materialise(st[1]))
                continue
            a, _, flag = pl.computeAction(*[np.array(x, dtype=float) for x in st])
            actions.append(np.array(a, dtype=float).tolist())
        golden[name] = {"robot": robot, "config": cfg, "setup": [[m, a] for m, a in setup], "steps": [list(s) for s in steps],
                        "N": pl._solver.N, "nvar": pl._solver.nvar, "calls": pl._solver.calls, "actions": actions}
        print(name, "calls", len(pl._solver.calls), "npar*N", len(pl._solver.calls[0]["all_parameters"]))
        if name == "pointRobot":
            # A dynamic obstacle (position, velocity, acceleration) riding behind the state (mpcPlanner.py:243-244) cannot
            # be pinned: the reference's updateDynamicObstacles calls self.dt(), which the class does not define (:158).
            try:
                pl.computeAction(np.zeros(3), np.zeros(3), np.arange(9.0))
                golden["updateDynamicObstacles"] = "ran"
            except AttributeError as e:
                golden["updateDynamicObstacles"] = "AttributeError: " + str(e)
            print("updateDynamicObstacles in the reference:", golden["updateDynamicObstacles"])
    with open(os.path.join(ROOT, "tests", "golden", "planner_reference.json"), "w") as f:
        json.dump(golden, f)


if __name__ == "__main__":
    main()
