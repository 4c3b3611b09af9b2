"""CPU-twin robustness/iteration statistics over the four workloads (algorithm experiments)."""
import os, sys, time, yaml, numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from robot_mpcs_b200 import scenarios as sc
from oracle.build_twin import Twin, build_twin
S = os.path.join(ROOT, 'robot_mpcs_b200', 'solvers')
W = [("pointRobot_n3_005_H20_noSlack", sc.point_robot_sweep, 2048), ("boxer_n3_01_H30_noSlack", sc.boxer_radial, 1024),
     ("panda_n7_005_H20_noSlack", sc.panda, 256), ("boxer_n3_01_H40_noSlack", sc.boxer_supermarket, 512)]
opts = None
if len(sys.argv) > 1:
    opts = np.array([float(x) for x in sys.argv[1].split(',')])
for sdir, gen, B in W:
    pm = yaml.safe_load(open(f'{S}/{sdir}/paramMap.yaml'))
    s = gen(pm, B)
    build_twin(f'{S}/{sdir}')
    tw = Twin(f'{S}/{sdir}')
    t = time.time(); r = tw.solve(s['xinit'], s['x0'], s['params'], opts=opts); dt = time.time() - t
    ef = r['exitflag']; it = r['iters']
    u, c = np.unique(ef, return_counts=True)
    print(f"{sdir:34s} B={B} {dt:5.2f}s {B/dt:7.0f}/s conv {np.mean(ef==1):.4f} flags {dict(zip(u.tolist(), c.tolist()))} iters mean {it.mean():.1f} p50 {np.median(it):.0f} p99 {np.quantile(it,0.99):.0f} max {it.max()}  bad {np.where(ef!=1)[0][:8].tolist()}")
