"""Generates tests/golden/local_goal_reference.json by running the REFERENCE's own GlobalPlanner.get_local_goal
(/root/reference/robotmpcs/global_planner/globalPlanner.py:174-189) on seeded paths and robot tracks.  Runs only in the
build container (the reference tree does not travel).  matplotlib and the reference's a_star / gridmap modules (pypng,
matplotlib file I/O — SURVEY.md §2 row 14, out of scope) are satisfied by empty stubs: get_local_goal touches none.

    python tests/golden/make_local_goal_golden.py
"""
import importlib.util
import json
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/robotmpcs/global_planner/globalPlanner.py"


def load_reference_class():
    for name in ("matplotlib", "matplotlib.pyplot", "robotmpcs", "robotmpcs.global_planner",
                 "robotmpcs.global_planner.a_star", "robotmpcs.global_planner.gridmap"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.modules["robotmpcs.global_planner.a_star"].a_star = None
    sys.modules["robotmpcs.global_planner.gridmap"].OccupancyGridMap = None
    spec = importlib.util.spec_from_file_location("ref_global_planner", REF)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.GlobalPlanner


def make_case(rng, L, T, threshold, speed):
    """A wandering path of L waypoints and a robot that chases the current local goal with noise (so that it comes
    within the threshold of some waypoints, overshoots others, and sometimes idles)."""
    path = np.cumsum(rng.uniform(-1.0, 1.0, (L, 2)) + np.array([0.8, 0.0]), axis=0)
    return path, threshold, speed


def main():
    GP = load_reference_class()
    rng = np.random.default_rng(0)
    cases = []
    for L, T, thr, speed in [(12, 80, 1.3, 0.35), (5, 40, 1.3, 0.5), (1, 5, 1.3, 0.3), (2, 10, 0.4, 0.2), (30, 120, 0.8, 0.6),
                             (8, 30, 1.3, 0.0)]:
        path, thr, speed = make_case(rng, L, T, thr, speed)
        gp = GP(dim_pixels=np.array([10, 10]), limits_low=np.array([0.0, 0.0]), limits_high=np.array([1.0, 1.0]),
                BOOL_PLOTTING=False, threshold_local_goal=thr)
        pos = path[0] + rng.uniform(-0.5, 0.5, 2)
        positions, goals, idxs = [], [], []
        for t in range(T):
            g = gp.get_local_goal(pos, [list(p) for p in path])
            positions.append(pos.tolist())
            goals.append([float(g[0]), float(g[1])])
            idxs.append(int(gp.idx_local))
            d = np.asarray(g) - pos
            n = np.linalg.norm(d)
            pos = pos + (speed * d / n if n > 1e-9 else 0.0) + rng.normal(0, 0.05, 2)
        cases.append({"path": path.tolist(), "threshold": thr, "positions": positions, "local_goal": goals, "idx_local": idxs})
    out = os.path.join(HERE, "local_goal_reference.json")
    json.dump({"source": REF + ":174-189", "cases": cases}, open(out, "w"))
    print("wrote", out, [c["idx_local"][-1] for c in cases], [len(c["path"]) for c in cases])


if __name__ == "__main__":
    main()
