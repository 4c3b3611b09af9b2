"""Solver object over the C ABI of a generated solver directory (include/rmpc.h).

Plays the role of `forcespro.nlp.Solver` for the planner:
    reference                                      here
    Solver.from_directory(dir)   mpcPlanner.py:73   Solver.from_directory(dir)   (dlopen lib<name>.so via ctypes)
    solver.solve(problem)        mpcPlanner.py:262  solver.solve(problem) -> (output{'x01'..}, exitflag, info)   B = 1
                                                    solver.solve_batch(xinit, x0, params)                      B >= 1
There is no CPU fallback: a missing library, a missing CUDA device or a failing kernel raises.
"""
from __future__ import annotations

import ctypes
import os
from typing import Optional

import numpy as np

try:  # torch is only the batch carrier (device memory + streams)
    import torch
except Exception:  # pragma: no cover
    torch = None


class RmpcDims(ctypes.Structure):
    _fields_ = [(n, ctypes.c_int) for n in ("N", "nx", "nu", "nvar", "npar", "nh")] + [("dt", ctypes.c_double)]


class RmpcOptions(ctypes.Structure):
    _fields_ = [("max_iter", ctypes.c_int), ("hessian", ctypes.c_int), ("precision", ctypes.c_int),
                ("control_mode", ctypes.c_int), ("tol_stat", ctypes.c_double), ("tol_eq", ctypes.c_double),
                ("tol_ineq", ctypes.c_double), ("tol_comp", ctypes.c_double), ("mu0", ctypes.c_double),
                ("mu_min", ctypes.c_double), ("reg", ctypes.c_double), ("creep_trigger", ctypes.c_int),
                ("stall_window", ctypes.c_int)]


class RmpcInfo(ctypes.Structure):
    _fields_ = [("exitflag", ctypes.c_void_p), ("iters", ctypes.c_void_p), ("obj", ctypes.c_void_p),
                ("res_stat", ctypes.c_void_p), ("res_eq", ctypes.c_void_p), ("res_ineq", ctypes.c_void_p),
                ("res_comp", ctypes.c_void_p)]


EXPORTS = ["rmpc_model_name", "rmpc_get_dims", "rmpc_default_options", "rmpc_last_error", "rmpc_create", "rmpc_destroy",
           "rmpc_set_options", "rmpc_get_options", "rmpc_solve_batch", "rmpc_solve_batch_strided", "rmpc_solve_batch_host", "rmpc_solve_batch_host_async", "rmpc_wait", "rmpc_fsd_constraints", "rmpc_shift_horizon",
           "rmpc_set_params", "rmpc_plant_step", "rmpc_set_x0", "rmpc_predict_obstacles", "rmpc_solve_batch_static", "rmpc_static_max_batch", "rmpc_launch_count",
           "rmpc_exchange_region_bytes", "rmpc_ipc_alloc", "rmpc_ipc_open", "rmpc_ipc_close", "rmpc_ipc_free", "rmpc_exchange_create",
           "rmpc_exchange_destroy", "rmpc_exchange_publish", "rmpc_exchange_collect", "rmpc_exchange_join", "rmpc_exchange_status"]


class SolverError(RuntimeError):
    pass


def load_library(solver_dir: str, lib_dir: Optional[str] = None) -> ctypes.CDLL:
    name = os.path.basename(os.path.normpath(solver_dir))
    path = os.path.join(lib_dir or solver_dir, f"lib{name}.so")
    if lib_dir and not os.path.isfile(path):
        path = os.path.join(solver_dir, f"lib{name}.so")
    if not os.path.isfile(path):
        raise SolverError(f"{path} not found: generate the solver first (MpcModel.generateSolver); there is no CPU fallback")
    lib = ctypes.CDLL(path)
    for sym in EXPORTS:
        getattr(lib, sym)  # AttributeError if the ABI is incomplete
    lib.rmpc_model_name.restype = ctypes.c_char_p
    lib.rmpc_last_error.restype = ctypes.c_char_p
    lib.rmpc_create.restype = ctypes.c_void_p
    lib.rmpc_create.argtypes = [ctypes.c_int, ctypes.c_int]
    lib.rmpc_destroy.argtypes = [ctypes.c_void_p]
    lib.rmpc_set_options.argtypes = [ctypes.c_void_p, ctypes.POINTER(RmpcOptions)]
    lib.rmpc_get_options.argtypes = [ctypes.c_void_p, ctypes.POINTER(RmpcOptions)]
    lib.rmpc_solve_batch.argtypes = [ctypes.c_void_p, ctypes.c_int] + [ctypes.c_void_p] * 5 + [ctypes.POINTER(RmpcInfo), ctypes.c_void_p]
    lib.rmpc_solve_batch_strided.argtypes = ([ctypes.c_void_p, ctypes.c_int] + [ctypes.c_void_p] * 3 + [ctypes.c_long, ctypes.c_long]
                                             + [ctypes.c_void_p] * 2 + [ctypes.POINTER(RmpcInfo), ctypes.c_void_p])
    lib.rmpc_solve_batch_host.argtypes = [ctypes.c_void_p, ctypes.c_int] + [ctypes.c_void_p] * 5 + [ctypes.POINTER(RmpcInfo)]
    lib.rmpc_solve_batch_host_async.argtypes = lib.rmpc_solve_batch_host.argtypes
    lib.rmpc_wait.argtypes = [ctypes.c_void_p]
    lib.rmpc_shift_horizon.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
    lib.rmpc_set_params.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.POINTER(ctypes.c_int), ctypes.c_int,
                                    ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]
    lib.rmpc_fsd_constraints.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_long,
                                         ctypes.c_long, ctypes.c_double, ctypes.c_double, ctypes.c_double, ctypes.c_int,
                                         ctypes.c_void_p, ctypes.c_void_p, ctypes.POINTER(ctypes.c_int), ctypes.c_void_p]
    lib.rmpc_set_x0.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]
    lib.rmpc_plant_step.argtypes = [ctypes.c_void_p, ctypes.c_int] + [ctypes.c_void_p] * 5
    lib.rmpc_predict_obstacles.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                           ctypes.c_double, ctypes.c_void_p, ctypes.c_void_p]
    lib.rmpc_solve_batch_static.argtypes = ([ctypes.c_int] + [ctypes.c_void_p] * 5 + [ctypes.POINTER(RmpcInfo), ctypes.POINTER(RmpcOptions), ctypes.c_void_p])
    lib.rmpc_launch_count.restype = ctypes.c_long
    lib.rmpc_exchange_region_bytes.restype = ctypes.c_size_t
    lib.rmpc_exchange_region_bytes.argtypes = [ctypes.c_int] * 3
    lib.rmpc_ipc_alloc.argtypes = [ctypes.c_int, ctypes.c_size_t, ctypes.POINTER(ctypes.c_void_p), ctypes.c_char_p]
    lib.rmpc_ipc_open.argtypes = [ctypes.c_int, ctypes.c_char_p, ctypes.POINTER(ctypes.c_void_p)]
    lib.rmpc_ipc_close.argtypes = [ctypes.c_int, ctypes.c_void_p]
    lib.rmpc_ipc_free.argtypes = [ctypes.c_int, ctypes.c_void_p]
    lib.rmpc_exchange_create.restype = ctypes.c_void_p
    lib.rmpc_exchange_create.argtypes = [ctypes.c_int] * 5 + [ctypes.POINTER(ctypes.c_void_p)]
    lib.rmpc_exchange_destroy.argtypes = [ctypes.c_void_p]
    lib.rmpc_exchange_publish.restype = ctypes.c_long
    lib.rmpc_exchange_publish.argtypes = [ctypes.c_void_p, ctypes.c_int] + [ctypes.c_void_p] * 4
    lib.rmpc_exchange_collect.argtypes = [ctypes.c_void_p, ctypes.c_long] + [ctypes.c_void_p] * 4
    lib.rmpc_exchange_join.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
    lib.rmpc_exchange_status.argtypes = [ctypes.c_void_p]
    return lib


class Solver:
    def __init__(self, solver_dir: str, device: int = 0, max_batch: int = 1, lib_dir: Optional[str] = None):
        self.dir = solver_dir
        self.lib = load_library(solver_dir, lib_dir)
        d = RmpcDims()
        self.lib.rmpc_get_dims(ctypes.byref(d))
        self.N, self.nx, self.nu, self.nvar, self.npar, self.nh, self.dt = d.N, d.nx, d.nu, d.nvar, d.npar, d.nh, d.dt
        self.device = device
        self.max_batch = 0
        self._h = None
        self._opts = RmpcOptions()
        self.lib.rmpc_default_options(ctypes.byref(self._opts))
        self._ensure(max_batch)

    # -- reference-shaped constructor (mpcPlanner.py:73)
    @classmethod
    def from_directory(cls, solver_dir: str, device: int = 0, max_batch: int = 1, lib_dir: Optional[str] = None) -> "Solver":
        return cls(solver_dir, device=device, max_batch=max_batch, lib_dir=lib_dir)

    def _err(self) -> str:
        return self.lib.rmpc_last_error().decode()

    def _ensure(self, B: int):
        if self._h is not None and B <= self.max_batch:
            return
        if self._h is not None:
            self.lib.rmpc_destroy(self._h)
            self._h = None
        cap = max(B, 1)
        h = self.lib.rmpc_create(self.device, cap)
        if not h:
            raise SolverError("rmpc_create failed: " + self._err())
        self._h, self.max_batch = h, cap
        self.lib.rmpc_set_options(self._h, ctypes.byref(self._opts))

    def set_options(self, **kw):
        for k, v in kw.items():
            if not hasattr(self._opts, k):
                raise KeyError(k)
            setattr(self._opts, k, v)
        if self.lib.rmpc_set_options(self._h, ctypes.byref(self._opts)) != 0:
            msg = self._err()
            self.lib.rmpc_get_options(self._h, ctypes.byref(self._opts))  # keep the mirror equal to what the handle holds
            raise SolverError(msg)

    def options(self) -> dict:
        return {n: getattr(self._opts, n) for n, _ in RmpcOptions._fields_}

    def launch_count(self) -> int:
        return int(self.lib.rmpc_launch_count())

    def __del__(self):
        try:
            if self._h is not None:
                self.lib.rmpc_destroy(self._h)
        except Exception:
            pass

    # -- batched solve -----------------------------------------------------------------------
    def solve_batch(self, xinit, x0, params, want_z=True, want_info=True, block=True):
        """xinit[B,nx], x0[B,N,nvar], params[B,N,npar] float64.  CUDA torch tensors -> device path (asynchronous on
        the current stream, returns CUDA tensors); numpy arrays / CPU tensors -> host path (returns numpy).
        Device path only: params[B,npar] (two dimensions) is the compressed form for parameters that do not vary over
        the horizon (rmpc_solve_batch_strided with stage stride 0)."""
        if torch is not None and isinstance(xinit, torch.Tensor) and xinit.is_cuda:
            return self._solve_device(xinit, x0, params, want_z, want_info)
        return self._solve_host(np.asarray(xinit), np.asarray(x0), np.asarray(params), want_z, want_info, block)

    def _solve_device(self, xinit, x0, params, want_z, want_info):
        B = xinit.shape[0]
        self._ensure(B)
        f64 = torch.float64
        xinit = xinit.to(f64).contiguous().view(B, self.nx)
        x0 = x0.to(f64).contiguous().view(B, self.N * self.nvar)
        if params.dim() == 2 and params.shape[1] == self.npar and self.N > 1:
            params = params.to(f64).contiguous()
            p_bs, p_ss = self.npar, 0
        else:
            params = params.to(f64).contiguous().view(B, self.N * self.npar)
            p_bs, p_ss = self.N * self.npar, self.npar
        dev = xinit.device
        z = torch.empty(B, self.N, self.nvar, dtype=f64, device=dev) if want_z else None
        act = torch.empty(B, self.nu, dtype=f64, device=dev)
        out = {"action": act, "z": z}
        info = RmpcInfo()
        if want_info:
            out["exitflag"] = torch.full((B,), -1, dtype=torch.int32, device=dev)   # -1 = never written (an internal error, include/rmpc.h)
            out["iters"] = torch.empty(B, dtype=torch.int32, device=dev)
            for k in ("obj", "res_stat", "res_eq", "res_ineq", "res_comp"):
                out[k] = torch.empty(B, dtype=f64, device=dev)
            for k, _ in RmpcInfo._fields_:
                setattr(info, k, out[k].data_ptr())
        stream = torch.cuda.current_stream(dev).cuda_stream
        rc = self.lib.rmpc_solve_batch_strided(self._h, B, xinit.data_ptr(), x0.data_ptr(), params.data_ptr(), p_bs, p_ss,
                                               z.data_ptr() if want_z else None, act.data_ptr(),
                                               ctypes.byref(info) if want_info else None, ctypes.c_void_p(stream))
        if rc != 0:
            raise SolverError("rmpc_solve_batch failed: " + self._err())
        return out

    def wait(self):
        """Block until the last `solve_batch(..., block=False)` on host buffers has delivered its results."""
        if self.lib.rmpc_wait(self._h) != 0:
            raise SolverError("rmpc_wait failed: " + self._err())

    def _solve_host(self, xinit, x0, params, want_z, want_info, block=True):
        xinit = np.ascontiguousarray(xinit, dtype=np.float64).reshape(-1, self.nx)
        B = xinit.shape[0]
        self._ensure(B)
        x0 = np.ascontiguousarray(x0, dtype=np.float64).reshape(B, self.N, self.nvar)
        params = np.ascontiguousarray(params, dtype=np.float64).reshape(B, self.N, self.npar)
        z = np.empty((B, self.N, self.nvar)) if want_z else None
        act = np.empty((B, self.nu))
        out = {"action": act, "z": z}
        info = RmpcInfo()
        if want_info:
            out["exitflag"] = np.empty(B, dtype=np.int32)
            out["iters"] = np.empty(B, dtype=np.int32)
            for k in ("obj", "res_stat", "res_eq", "res_ineq", "res_comp"):
                out[k] = np.empty(B)
            for k, _ in RmpcInfo._fields_:
                setattr(info, k, out[k].ctypes.data)
        self._keep = (xinit, x0, params)  # the copies are asynchronous: keep the (possibly converted) inputs alive
        fn = self.lib.rmpc_solve_batch_host if block else self.lib.rmpc_solve_batch_host_async
        rc = fn(self._h, B, xinit.ctypes.data, x0.ctypes.data, params.ctypes.data,
                                            z.ctypes.data if want_z else None, act.ctypes.data,
                                            ctypes.byref(info) if want_info else None)
        if rc != 0:
            raise SolverError("rmpc_solve_batch_host failed: " + self._err())
        return out

    # -- reference-shaped single solve (mpcPlanner.py:246-262) ------------------------------------
    def solve(self, problem: dict):
        """problem = {'xinit': [nx], 'x0': [N*nvar], 'all_parameters': [N*npar]} -> (output, exitflag, info) with
        output['x01'..'xNN'] (key width as mpcPlanner.py:265-273)."""
        r = self._solve_host(np.asarray(problem["xinit"]).reshape(1, -1), np.asarray(problem["x0"]).reshape(1, -1),
                             np.asarray(problem["all_parameters"]).reshape(1, -1), True, True)
        width = len(str(self.N))
        output = {"x" + str(k + 1).zfill(width): r["z"][0, k].copy() for k in range(self.N)}
        info = {k: (int(r[k][0]) if k == "iters" else float(r[k][0])) for k in ("iters", "obj", "res_stat", "res_eq", "res_ineq", "res_comp")}
        return output, int(r["exitflag"][0]), info

    # -- warm start / parameter kernels --------------------------------------------------------
    def shift_horizon(self, z, x0=None):
        """previous-plan warm start on the device (mpcPlanner.py:215-226)."""
        B = z.shape[0]
        self._ensure(B)
        if x0 is None:
            x0 = torch.empty_like(z)
        stream = torch.cuda.current_stream(z.device).cuda_stream
        if self.lib.rmpc_shift_horizon(self._h, B, z.data_ptr(), x0.data_ptr(), ctypes.c_void_p(stream)) != 0:
            raise SolverError(self._err())
        return x0

    def set_x0(self, xinit, x0):
        """x0[B,N,nvar] (CUDA): state part of every stage <- xinit[B,nx] (mpcPlanner.py:228-232)."""
        B = x0.shape[0]
        self._ensure(B)
        xinit = xinit.to(torch.float64).contiguous()
        stream = torch.cuda.current_stream(x0.device).cuda_stream
        if self.lib.rmpc_set_x0(self._h, B, xinit.data_ptr(), x0.data_ptr(), ctypes.c_void_p(stream)) != 0:
            raise SolverError(self._err())
        return x0

    def plant_step(self, x, u, exitflag=None, out=None):
        """x+[B,nx] of the model the solver was generated from (the plant of a headless closed loop); problems with
        exitflag < 0 are stepped with u = 0.  CUDA float64 tensors; `out` may be `x`."""
        B = x.shape[0]
        self._ensure(B)
        x = x.to(torch.float64).contiguous()
        u = u.to(torch.float64).contiguous()
        if out is None:
            out = torch.empty_like(x)
        ef = None
        if exitflag is not None:
            ef = exitflag.to(torch.int32).contiguous()
        stream = torch.cuda.current_stream(x.device).cuda_stream
        if self.lib.rmpc_plant_step(self._h, B, x.data_ptr(), u.data_ptr(), ef.data_ptr() if ef is not None else None,
                                    out.data_ptr(), ctypes.c_void_p(stream)) != 0:
            raise SolverError(self._err())
        return out

    def set_params(self, params, idx, values):
        """params[B,N,npar] (CUDA, float64) <- values[B,len(idx)] or values[B,N,len(idx)] at paramMap indices idx."""
        B = params.shape[0]
        self._ensure(B)
        idx_arr = (ctypes.c_int * len(idx))(*[int(i) for i in idx])
        values = values.to(torch.float64).contiguous()
        sv = 1 if values.dim() == 3 else 0
        stream = torch.cuda.current_stream(params.device).cuda_stream
        if self.lib.rmpc_set_params(self._h, B, params.data_ptr(), idx_arr, len(idx), values.data_ptr(), sv,
                                    ctypes.c_void_p(stream)) != 0:
            raise SolverError(self._err())
        return params

    def predict_obstacles(self, params, obst_state, n_slots, p_obst, radius):
        """params[B,N,npar] (CUDA) <- constant-acceleration prediction of the moving obstacles obst_state[B,n_dyn,9]
        (mpcPlanner.py:144-161) in the `obst` slots starting at parameter index p_obst."""
        B = params.shape[0]
        self._ensure(B)
        n_dyn = 0 if obst_state is None else obst_state.shape[1]
        st = obst_state.to(torch.float64).contiguous() if n_dyn else None
        stream = torch.cuda.current_stream(params.device).cuda_stream
        if self.lib.rmpc_predict_obstacles(self._h, B, st.data_ptr() if n_dyn else None, n_dyn, n_slots, p_obst, float(radius),
                                           params.data_ptr(), ctypes.c_void_p(stream)) != 0:
            raise SolverError(self._err())
        return params

    def fsd_constraints(self, points, ref, K, max_radius=5.0, params=None, idx=None, want_out=True, lidar_offset=0.4, lidar_z=0.02):
        """Batched free-space decomposition on the device (robotmpcs/utils/free_space_decomposition.py as driven by
        examples/boxer_example.py:132-148,191-203).  points[B,P,3] (CUDA, float64); ref[B,N,>=3] (e.g. the previous plan
        z[B,N,nvar]) or ref[B,>=3] (one state per problem, used for every stage).  Returns [B,N,K,4] and/or writes the planes
        into params[B,N,npar] at the `lin_constrs_c` indices idx[c]."""
        B, P = points.shape[0], points.shape[1]
        points = points.to(torch.float64).contiguous()
        ref = ref.to(torch.float64)
        if ref.stride(-1) != 1:
            ref = ref.contiguous()
        if ref.dim() == 2:
            bs, ss = ref.stride(0), 0
        else:
            if ref.shape[1] != self.N:
                raise SolverError(f"ref has {ref.shape[1]} stages, the solver {self.N}")
            bs, ss = ref.stride(0), ref.stride(1)
        out = torch.empty(B, self.N, K, 4, dtype=torch.float64, device=points.device) if want_out else None
        idx_arr = (ctypes.c_int * K)(*[int(i) for i in idx]) if idx is not None else None
        stream = torch.cuda.current_stream(points.device).cuda_stream
        rc = self.lib.rmpc_fsd_constraints(self._h, B, points.data_ptr(), P, ref.data_ptr(), bs, ss, max_radius, lidar_offset, lidar_z,
                                           K, out.data_ptr() if want_out else None, params.data_ptr() if params is not None else None,
                                           idx_arr, ctypes.c_void_p(stream))
        if rc != 0:
            raise SolverError("rmpc_fsd_constraints failed: " + self._err())
        return out
