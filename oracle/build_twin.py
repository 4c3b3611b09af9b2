"""ORACLE build helper (test/bench infrastructure): compiles oracle/cpu_twin.cpp against a generated solver
directory's model_traits.h -> <solver_dir>/libtwin_<name>.so (host code only, g++ -fopenmp)."""
import ctypes
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def build_twin(solver_dir: str, force=False) -> str:
    name = os.path.basename(os.path.normpath(solver_dir))
    so = os.path.join(solver_dir, f"libtwin_{name}.so")
    src = os.path.join(ROOT, "oracle", "cpu_twin.cpp")
    deps = [src, os.path.join(solver_dir, "model_traits.h"), os.path.join(ROOT, "robot_mpcs_b200", "csrc", "rmpc_ipm.cuh"),
            os.path.join(ROOT, "robot_mpcs_b200", "csrc", "rmpc_model.cuh")]
    if not force and os.path.isfile(so) and all(os.path.getmtime(so) >= os.path.getmtime(d) for d in deps):
        return so
    cmd = ["g++", "-O3", "-march=native", "-fopenmp", "-std=c++17", "-shared", "-fPIC", "-x", "c++",
           "-I", solver_dir, "-I", os.path.join(ROOT, "robot_mpcs_b200", "csrc"), src, "-o", so]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("g++ failed:\n" + res.stdout + res.stderr)
    return so


class Twin:
    def __init__(self, solver_dir: str):
        self.lib = ctypes.CDLL(build_twin(solver_dir))
        d = (ctypes.c_int * 6)()
        self.lib.twin_dims(d)
        self.N, self.nx, self.nu, self.nvar, self.npar, self.nh = list(d)
        self.lib.twin_max_threads.restype = ctypes.c_int

    @staticmethod
    def _p(a):
        return a.ctypes.data_as(ctypes.c_void_p)

    def solve(self, xinit, x0, params, opts=None, nthreads=0, precision=0):
        xinit = np.ascontiguousarray(xinit, dtype=np.float64).reshape(-1, self.nx)
        B = xinit.shape[0]
        x0 = np.ascontiguousarray(x0, dtype=np.float64).reshape(B, self.N, self.nvar)
        params = np.ascontiguousarray(params, dtype=np.float64).reshape(B, self.N, self.npar)
        z = np.zeros((B, self.N, self.nvar))
        ef = np.zeros(B, dtype=np.int32)
        it = np.zeros(B, dtype=np.int32)
        res = np.zeros((B, 6))
        o = None
        if opts is not None:
            o = np.zeros(32)                       # trailing entries the caller did not give: 0
            o[: len(opts)] = np.asarray(opts, dtype=np.float64)
        self.lib.twin_solve_batch(B, self._p(xinit), self._p(x0), self._p(params), self._p(z), self._p(ef), self._p(it),
                                  self._p(res), self._p(o) if o is not None else None, nthreads, precision)
        return {"z": z, "exitflag": ef, "iters": it, "obj": res[:, 0], "res_stat": res[:, 1], "res_eq": res[:, 2],
                "res_ineq": res[:, 3], "res_comp": res[:, 4], "mu": res[:, 5]}

    def eval_stage(self, z, p):
        """(J(z, p), h(z, p)) of one stage through csrc/rmpc_model.cuh:stage_values — the model code the kernel compiles."""
        z = np.ascontiguousarray(z, dtype=np.float64)
        p = np.ascontiguousarray(p, dtype=np.float64)
        f = ctypes.c_double(0.0)
        h = np.zeros(self.nh)
        self.lib.twin_stage_values(self._p(z), self._p(p), ctypes.byref(f), self._p(h))
        return f.value, h
