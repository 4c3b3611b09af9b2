"""Generates the golden fixtures tests/golden/*.json with the ORACLE (oracle/nlp.py + oracle/solve.py, scipy SLSQP in
float64).  Run from the repo root:  python tests/golden/make_golden.py [name ...]

The reference ships no golden vectors and its solver stack cannot run here, so these are the pins.
For every scenario the oracle is started (a) from the planner's own initial guess x0 and (b) from the candidate
produced by the CPU twin (to make sure both sit in the same basin of this non-convex NLP); the fixture stores
(b) polished by SLSQP, its solver-independent KKT certificate, and whether (a) reached the same point.
"""
import json
import os
import sys
import time

import numpy as np
import yaml

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle.nlp import load_model  # noqa: E402
from oracle.solve import solve_slsqp, kkt_certificate  # noqa: E402
from oracle.build_twin import Twin  # noqa: E402
from robot_mpcs_b200 import scenarios as sc  # noqa: E402

CFG = os.path.join(ROOT, "robot_mpcs_b200", "config")
SOLVERS = os.path.join(ROOT, "robot_mpcs_b200", "solvers")

CASES = {
    # fixture name: (config, solver dir, scenario generator, kwargs, indices, also solve from x0?)
    "pointRobot_example": ("pointRobotMpc.yaml", "pointRobot_n3_005_H20_noSlack", sc.point_robot_example, {}, [0], True),
    "pointRobot_sweep": ("pointRobotMpc.yaml", "pointRobot_n3_005_H20_noSlack", sc.point_robot_sweep, {"B": 16}, [0, 1, 2, 3, 4, 5], False),
    "boxer_radial": ("boxerRadialMpc.yaml", "boxer_n3_01_H30_noSlack", sc.boxer_radial, {"B": 16}, [1, 2, 4], False),
    "boxer_supermarket": ("boxerSupermarketMpc.yaml", "boxer_n3_01_H40_noSlack", sc.boxer_supermarket, {"B": 8}, [0, 1], False),
    "panda_example": ("pandaMpc.yaml", "panda_n7_005_H20_noSlack", sc.panda, {"B": 1, "jitter": False}, [0], True),
    "panda_jitter": ("pandaMpc.yaml", "panda_n7_005_H20_noSlack", sc.panda, {"B": 8}, [0, 1], False),
}


def main(names):
    for name in names:
        cfg, sdir, gen, kw, idxs, from_x0 = CASES[name]
        model = load_model(os.path.join(CFG, cfg))
        pm = yaml.safe_load(open(os.path.join(SOLVERS, sdir, "paramMap.yaml")))
        assert pm == model.paramMap, "product and oracle paramMaps differ"
        s = gen(pm, **kw)
        tw = Twin(os.path.join(SOLVERS, sdir))
        cand = tw.solve(s["xinit"], s["x0"], s["params"])
        out = {"config": cfg, "solver": sdir, "generator": gen.__name__, "kwargs": kw, "cases": []}
        for b in idxs:
            t = time.time()
            xi, x0, P = s["xinit"][b], s["x0"][b], s["params"][b].reshape(-1)
            r = solve_slsqp(model, xi, cand["z"][b], P)
            cert = kkt_certificate(model, xi, P, r["Z"])
            case = {"index": b, "Z": r["Z"].tolist(), "obj": cert["obj"], "kkt": cert, "twin_exitflag": int(cand["exitflag"][b]),
                    "twin_vs_oracle": float(np.abs(cand["z"][b] - r["Z"]).max())}
            if from_x0:
                r0 = solve_slsqp(model, xi, x0, P)
                case["from_x0_max_abs_diff"] = float(np.abs(r0["Z"] - r["Z"]).max())
                case["from_x0_obj"] = r0["obj"]
            print(name, b, "obj %.9g" % cert["obj"], "kkt", {k: ("%.2e" % v if isinstance(v, float) else v) for k, v in cert.items()},
                  "twin-vs-oracle %.2e" % case["twin_vs_oracle"], "from_x0", case.get("from_x0_max_abs_diff"), "%.0fs" % (time.time() - t), flush=True)
            out["cases"].append(case)
        if len(idxs) == 1 and name.endswith("example"):
            out.update({"Z": out["cases"][0]["Z"], "obj": out["cases"][0]["obj"]})
        with open(os.path.join(ROOT, "tests", "golden", name + ".json"), "w") as f:
            json.dump(out, f)


if __name__ == "__main__":
    main(sys.argv[1:] or list(CASES))
