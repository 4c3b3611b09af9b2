"""Host twin: precision 1 (float32 iterations, then float64) against precision 0 on the four workloads."""
import os, sys, yaml, numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
from robot_mpcs_b200 import scenarios as sc
from oracle.build_twin import Twin
S = os.path.join(ROOT, "robot_mpcs_b200", "solvers")
W = [("pointRobot_n3_005_H20_noSlack", sc.point_robot_sweep, 2048), ("boxer_n3_01_H30_noSlack", sc.boxer_radial, 1024),
     ("panda_n7_005_H20_noSlack", sc.panda, 256), ("boxer_n3_01_H40_noSlack", sc.boxer_supermarket, 512)]
for sd, gen, B in W:
    pm = yaml.safe_load(open(f"{S}/{sd}/paramMap.yaml")); s = gen(pm, B); tw = Twin(f"{S}/{sd}")
    ref = tw.solve(s["xinit"], s["x0"], s["params"])
    m = tw.solve(s["xinit"], s["x0"], s["params"], precision=1)
    ok = (ref["exitflag"] == 1) & (m["exitflag"] == 1)
    err = np.abs(m["z"][ok] - ref["z"][ok]).max(axis=(1, 2)) / np.abs(ref["z"][ok]).max(axis=(1, 2))
    print(f"{sd:32s} f64 it {ref['iters'].mean():5.1f} p99 {np.quantile(ref['iters'], .99):.0f} conv {np.mean(ref['exitflag'] == 1):.4f} | mixed it {m['iters'].mean():5.1f} "
          f"p99 {np.quantile(m['iters'], .99):.0f} conv {np.mean(m['exitflag'] == 1):.4f} flags {dict(zip(*np.unique(m['exitflag'], return_counts=True)))} | same basin {np.mean(err < 1e-5):.3f} med err {np.median(err):.1e}")
