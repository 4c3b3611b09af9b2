"""Basin agreement of the product's algorithm with the oracle's golden fixtures (tests/golden/*.json, every case solved by
SLSQP from the planner's own x0).  The NLP is non-convex, so two correct solvers started from the same x0 can stop in
different local minima; this script counts, per fixture, how often the product lands on the oracle's point (trajectory
within 1e-4 relative, objective within 1e-5 relative), and for the disagreements whose objective is lower.

  python scripts/basin_agreement.py            CPU twin (the same Ipm<> template the kernel compiles; runs anywhere)
  python scripts/basin_agreement.py --gpu      the CUDA path through the C ABI (needs a GPU)
"""
import argparse, json, os, sys
import numpy as np, yaml
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from robot_mpcs_b200 import scenarios as sc
S = os.path.join(ROOT, "robot_mpcs_b200", "solvers")
G = os.path.join(ROOT, "tests", "golden")
FIX = ["pointRobot_example", "pointRobot_sweep", "boxer_radial", "panda_example", "panda_jitter", "boxer_supermarket", "boxer_linear", "boxer_vel"]


def valid(case):
    k = case["kkt"]
    return k["eq"] < 1e-6 and k["ineq"] < 1e-6 and k["stat"] < 1e-5


def compare(g, z, ef, obj):
    rows = []
    for c in g["cases"]:
        b = c["index"]
        zo = np.array(c["Z"])
        rel = float(np.abs(z[b] - zo).max() / np.abs(zo).max())
        orel = float(abs(obj[b] - c["obj"]) / abs(c["obj"]))
        rows.append({"index": b, "oracle_valid": valid(c), "exitflag": int(ef[b]), "rel": rel, "obj_rel": orel, "obj": float(obj[b]), "oracle_obj": c["obj"],
                     "same": bool(ef[b] == 1 and rel < 1e-4 and orel < 1e-5)})
    return rows


def summarise(name, rows):
    ok = [r for r in rows if r["oracle_valid"]]
    conv = [r for r in ok if r["exitflag"] == 1]
    same = [r for r in conv if r["same"]]
    diff = [r for r in conv if not r["same"]]
    worse = [r for r in diff if r["obj"] > r["oracle_obj"] * (1 + 1e-9) + 1e-12]
    return {"fixture": name, "cases": len(rows), "oracle_valid": len(ok), "converged": len(conv), "same_basin": len(same),
            "other_basin": len(diff), "other_basin_worse_objective": len(worse), "other_basin_better_objective": len(diff) - len(worse),
            "agreement": len(same) / max(1, len(ok)), "median_rel_same": float(np.median([r["rel"] for r in same])) if same else None,
            "max_rel_same": float(max(r["rel"] for r in same)) if same else None,
            "other": [(r["index"], round(r["obj"], 6), round(r["oracle_obj"], 6)) for r in diff],
            "not_converged": [(r["index"], r["exitflag"]) for r in ok if r["exitflag"] != 1]}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpu", action="store_true")
    ap.add_argument("--json", default=None)
    a = ap.parse_args()
    out = []
    for name in FIX:
        g = json.load(open(os.path.join(G, name + ".json")))
        pm = yaml.safe_load(open(os.path.join(S, g["solver"], "paramMap.yaml")))
        s = getattr(sc, g["generator"])(pm, **g["kwargs"])
        if a.gpu:
            import torch
            from robot_mpcs_b200.planner.solver import Solver
            sv = Solver.from_directory(os.path.join(S, g["solver"]), device=0, max_batch=s["xinit"].shape[0])
            r = sv.solve_batch(*(torch.as_tensor(s[k], device="cuda:0") for k in ("xinit", "x0", "params")))
            z, ef, obj = r["z"].cpu().numpy(), r["exitflag"].cpu().numpy(), r["obj"].cpu().numpy()
        else:
            from oracle.build_twin import Twin
            r = Twin(os.path.join(S, g["solver"])).solve(s["xinit"], s["x0"], s["params"])
            z, ef, obj = r["z"], r["exitflag"], r["obj"]
        sm = summarise(name, compare(g, z, ef, obj))
        out.append(sm)
        print(f"{name:20s} valid {sm['oracle_valid']:3d}/{sm['cases']:3d} converged {sm['converged']:3d} same basin {sm['same_basin']:3d} ({sm['agreement']:.3f}) "
              f"other {sm['other_basin']:2d} (worse obj {sm['other_basin_worse_objective']}, better {sm['other_basin_better_objective']}) "
              f"rel med {sm['median_rel_same']} max {sm['max_rel_same']}  not converged {sm['not_converged']}", flush=True)
        for o in sm["other"]:
            print("      other basin: index %d  ours %.6f  oracle %.6f" % o)
    if a.json:
        json.dump(out, open(a.json, "w"), indent=1)


if __name__ == "__main__":
    main()
