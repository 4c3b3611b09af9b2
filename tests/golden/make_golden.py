"""Generates the golden fixtures tests/golden/*.json with the ORACLE (oracle/nlp.py + oracle/solve.py, scipy SLSQP in
float64).  Run from the repo root:  python tests/golden/make_golden.py [--workers W] [name ...]

The reference ships no golden vectors and its solver stack cannot run here, so these are the pins.
EVERY case is solved by SLSQP **from the planner's own initial guess x0** — what the reference hands its solver
(/root/reference/robotmpcs/planner/mpcPlanner.py:246-250,262: problem = {xinit, x0, all_parameters}) — and never from
anything the product's algorithm (CUDA kernel or its host twin) produced: the fixtures do not depend on the code under
test.  Stored per case: the oracle's solution Z, its objective, SLSQP's own status and the solver-independent KKT
certificate (oracle/solve.py:kkt_certificate); for the first `TC` cases of every fixture a second, unrelated solver
(scipy trust-constr, interior point) is started from the same x0 and its distance to the SLSQP point is recorded.
The NLP is non-convex (obstacles), so two correct solvers started from the same x0 may stop in different local minima;
which fixtures the product reproduces (basin agreement) is measured by tests/test_gpu_parity.py::test_golden and
scripts/basin_agreement.py, not assumed here.

Indices are contiguous (0 .. n-1 of the seeded batch); nothing is skipped.
"""
import argparse
import json
import os
import sys
import time

os.environ.setdefault("OMP_NUM_THREADS", "1")
os.environ.setdefault("OPENBLAS_NUM_THREADS", "1")
os.environ.setdefault("MKL_NUM_THREADS", "1")

import numpy as np
import yaml

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

CFG = os.path.join(ROOT, "robot_mpcs_b200", "config")
SOLVERS = os.path.join(ROOT, "robot_mpcs_b200", "solvers")
TC = 8   # cases per fixture cross-checked with trust-constr

CASES = {
    # fixture name: (config, solver dir, scenario generator, kwargs, number of cases)
    "pointRobot_example": ("pointRobotMpc.yaml", "pointRobot_n3_005_H20_noSlack", "point_robot_example", {}, 1),
    "pointRobot_sweep": ("pointRobotMpc.yaml", "pointRobot_n3_005_H20_noSlack", "point_robot_sweep", {"B": 64}, 64),
    "boxer_radial": ("boxerRadialMpc.yaml", "boxer_n3_01_H30_noSlack", "boxer_radial", {"B": 64}, 64),
    "boxer_supermarket": ("boxerSupermarketMpc.yaml", "boxer_n3_01_H40_noSlack", "boxer_supermarket", {"B": 64}, 64),
    "panda_example": ("pandaMpc.yaml", "panda_n7_005_H20_noSlack", "panda", {"B": 1, "jitter": False}, 1),
    "panda_jitter": ("pandaMpc.yaml", "panda_n7_005_H20_noSlack", "panda", {"B": 64}, 64),
    "boxer_linear": ("boxerMpc.yaml", "boxer_n3_01_H10_noSlack", "boxer_linear", {"B": 64}, 64),
    "boxer_vel": ("boxerVelMpc.yaml", "boxerVel_n3_01_H15_noSlack", "boxer_vel", {"B": 64}, 64),
}

_cache = {}


def _setup(name):
    if name not in _cache:
        import torch
        torch.set_num_threads(1)
        from oracle.nlp import load_model
        from robot_mpcs_b200 import scenarios as sc
        cfg, sdir, gen, kw, n = CASES[name]
        model = load_model(os.path.join(CFG, cfg))
        pm = yaml.safe_load(open(os.path.join(SOLVERS, sdir, "paramMap.yaml")))
        assert pm == model.paramMap, "product and oracle paramMaps differ"
        _cache[name] = (model, getattr(sc, gen)(pm, **kw))
    return _cache[name]


def solve_case(task):
    name, b = task
    from oracle.solve import solve_slsqp, kkt_certificate
    model, s = _setup(name)
    t = time.time()
    xi, x0, P = s["xinit"][b], s["x0"][b], s["params"][b].reshape(-1)
    r = solve_slsqp(model, xi, x0, P)
    cert = kkt_certificate(model, xi, P, r["Z"])
    case = {"index": b, "Z": r["Z"].tolist(), "obj": cert["obj"], "kkt": cert, "slsqp": {"success": r["success"], "nit": r["nit"], "message": str(r["message"])},
            "start": "x0", "seconds": time.time() - t}
    print(name, b, "obj %.9g" % cert["obj"], "stat %.1e eq %.1e ineq %.1e act %d" % (cert["stat"], cert["eq"], cert["ineq"], cert["n_active"]),
          r["message"], "%.0fs" % case["seconds"], flush=True)
    return name, case


def cross_check(task):
    """Second solver (scipy trust-constr: interior point, BFGS Hessian) from the same x0; minutes per case."""
    name, b, Z = task
    from oracle.solve import solve_trust_constr, kkt_certificate
    model, s = _setup(name)
    t = time.time()
    xi, x0, P = s["xinit"][b], s["x0"][b], s["params"][b].reshape(-1)
    tc = solve_trust_constr(model, xi, x0, P)
    ctc = kkt_certificate(model, xi, P, tc["Z"])
    out = {"max_abs_diff": float(np.abs(tc["Z"] - np.asarray(Z)).max()), "obj": ctc["obj"], "nit": int(tc["nit"]),
           "kkt": {k: ctc[k] for k in ("eq", "ineq", "stat")}, "seconds": time.time() - t}
    print(name, b, "trust-constr diff %.2e obj %.9g stat %.1e %.0fs" % (out["max_abs_diff"], out["obj"], ctc["stat"], out["seconds"]), flush=True)
    return name, b, out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("names", nargs="*")
    ap.add_argument("--workers", type=int, default=max(1, (os.cpu_count() or 2) - 2))
    ap.add_argument("--no-cross-check", action="store_true")
    args = ap.parse_args()
    names = args.names or list(CASES)
    tasks = [(n, b) for n in names for b in range(CASES[n][4])]
    from multiprocessing import Pool
    results = {n: [] for n in names}
    with Pool(args.workers) as pool:
        for n, case in pool.imap_unordered(solve_case, tasks, chunksize=1):
            results[n].append(case)
            if len(results[n]) == CASES[n][4]:
                results[n].sort(key=lambda c: c["index"])
                write(n, results[n])
        if args.no_cross_check:
            return
        tasks = [(n, b, results[n][b]["Z"]) for b in range(TC) for n in names if b < CASES[n][4]]
        left = {n: min(TC, CASES[n][4]) for n in names}
        for n, b, out in pool.imap_unordered(cross_check, tasks, chunksize=1):
            results[n][b]["trust_constr"] = out
            left[n] -= 1
            if left[n] == 0:
                write(n, results[n])


def write(name, cases):
    cfg, sdir, gen, kw, n = CASES[name]
    out = {"config": cfg, "solver": sdir, "generator": gen, "kwargs": kw, "start": "x0 (planner's initial guess), scipy SLSQP", "cases": cases}
    if n == 1:
        out.update({"Z": cases[0]["Z"], "obj": cases[0]["obj"]})
    with open(os.path.join(ROOT, "tests", "golden", name + ".json"), "w") as f:
        json.dump(out, f)
    cert = [c["kkt"] for c in cases]
    print(f"== {name}: {len(cases)} cases written; worst certificate stat {max(c['stat'] for c in cert):.1e} eq {max(c['eq'] for c in cert):.1e} "
          f"ineq {max(c['ineq'] for c in cert):.1e}", flush=True)


if __name__ == "__main__":
    main()
