"""Trace one problem through the CPU twin (RMPC_TRACE build)."""
import os, sys, subprocess, ctypes, yaml, numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from robot_mpcs_b200 import scenarios as sc
from oracle.build_twin import Twin
S = os.path.join(ROOT, 'robot_mpcs_b200', 'solvers')
W = {"pr": ("pointRobot_n3_005_H20_noSlack", sc.point_robot_sweep, 2048), "boxer": ("boxer_n3_01_H30_noSlack", sc.boxer_radial, int(os.environ.get("NB", 1024))),
     "panda": ("panda_n7_005_H20_noSlack", sc.panda, 256), "sm": ("boxer_n3_01_H40_noSlack", sc.boxer_supermarket, 512)}
sdir, gen, B = W[sys.argv[1]]
b = int(sys.argv[2])
opts = np.array([float(x) for x in sys.argv[3].split(',')]) if len(sys.argv) > 3 else None
so = f'/tmp/twin_trace_{sys.argv[1]}.so'
subprocess.check_call(["g++", "-O3", "-DRMPC_TRACE=" + os.environ.get("TRACE", "1"), "-march=native", "-fopenmp", "-std=c++17", "-shared", "-fPIC", "-x", "c++", "-I", f"{S}/{sdir}",
                       "-I", f"{ROOT}/robot_mpcs_b200/csrc", f"{ROOT}/oracle/cpu_twin.cpp", "-o", so])
pm = yaml.safe_load(open(f'{S}/{sdir}/paramMap.yaml'))
s = gen(pm, B)
tw = Twin(f'{S}/{sdir}')
tw.lib = ctypes.CDLL(so); tw.lib.twin_max_threads.restype = ctypes.c_int
r = tw.solve(s['xinit'][b:b+1], s['x0'][b:b+1], s['params'][b:b+1], opts=opts, nthreads=1)
print('exitflag', r['exitflag'], 'iters', r['iters'])
