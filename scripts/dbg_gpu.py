"""GPU-vs-twin divergence probe (debug helper, not a test)."""
import os, sys, yaml, numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from robot_mpcs_b200 import scenarios as sc
from robot_mpcs_b200.planner.solver import Solver
from oracle.build_twin import Twin
S = os.path.join(ROOT, 'robot_mpcs_b200', 'solvers', 'boxer_n3_01_H30_noSlack')
B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
pm = yaml.safe_load(open(S + '/paramMap.yaml'))
s = sc.boxer_radial(pm, B)
sol = Solver.from_directory(S, 0, B)
for hess in (1, 0):
    sol.set_options(hessian=hess)
    d = {k: torch.as_tensor(v, device='cuda:0') for k, v in s.items()}
    r = sol.solve_batch(d['xinit'], d['x0'], d['params'])
    ef = r['exitflag'].cpu().numpy(); it = r['iters'].cpu().numpy()
    opts = np.array([200, hess, 1e-6, 1e-8, 1e-8, 1e-8, 1.0, 1e-9, 0.0, 1e-2, 1e10, 0.99, 1e-2, 4.0, 0.25, 3])
    tw = Twin(S).solve(s['xinit'], s['x0'], s['params'], opts=opts)
    print('hessian', hess, 'gpu flags', np.unique(ef, return_counts=True), 'twin flags', np.unique(tw['exitflag'], return_counts=True))
    bad = np.where(ef != tw['exitflag'])[0]
    print(' mismatch idx', bad[:40])
    for b in bad[:8]:
        print('  b', b, 'gpu', ef[b], it[b], 'stat %.2e eq %.2e ineq %.2e comp %.2e' % tuple(float(r[k][b]) for k in ('res_stat', 'res_eq', 'res_ineq', 'res_comp')),
              '| twin', tw['exitflag'][b], tw['iters'][b], 'stat %.2e comp %.2e mu %.2e' % (tw['res_stat'][b], tw['res_comp'][b], tw['mu'][b]))
    # single-problem re-run of the mismatching ones
    for b in bad[:4]:
        r1 = sol.solve_batch(d['xinit'][b:b+1], d['x0'][b:b+1], d['params'][b:b+1])
        print('  alone b', b, int(r1['exitflag'][0]), int(r1['iters'][0]))
    same = (ef == 1) & (tw['exitflag'] == 1)
    print(' iters equal frac', (it[same] == tw['iters'][same]).mean(), 'max z diff', np.abs(r['z'].cpu().numpy()[same] - tw['z'][same]).max())
