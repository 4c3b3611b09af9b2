"""rmpc_solve_batch_static (include/rmpc.h): the batched solve whose scratch memory is a file-scope device array — no
handle, no cudaMalloc.  No counterpart in the reference.  Called through the
C ABI on two batches in a row (the second smaller, with infeasible starts, so that stale counts or list entries of the
first call would show) and compared with rmpc_solve_batch; a batch beyond the stated maximum is refused.
Collected last on purpose (file name)."""
import ctypes
import os

import numpy as np
import pytest
import yaml

from conftest import SOLVERS
from robot_mpcs_b200 import scenarios as sc

pytestmark = pytest.mark.gpu
SDIR = "boxer_n3_01_H30_noSlack"


def _call_static(lib, d, B, N, nvar, nu):
    import torch
    from robot_mpcs_b200.planner.solver import RmpcInfo
    dev = d["xinit"].device
    out = {"z": torch.empty(B, N, nvar, dtype=torch.float64, device=dev), "action": torch.empty(B, nu, dtype=torch.float64, device=dev),
           "exitflag": torch.full((B,), -1, dtype=torch.int32, device=dev), "iters": torch.zeros(B, dtype=torch.int32, device=dev)}
    for k in ("obj", "res_stat", "res_eq", "res_ineq", "res_comp"):
        out[k] = torch.empty(B, dtype=torch.float64, device=dev)
    info = RmpcInfo()
    for k, _ in RmpcInfo._fields_:
        setattr(info, k, out[k].data_ptr())
    stream = torch.cuda.current_stream(dev).cuda_stream
    rc = lib.rmpc_solve_batch_static(B, d["xinit"].data_ptr(), d["x0"].data_ptr(), d["params"].data_ptr(), out["z"].data_ptr(),
                                     out["action"].data_ptr(), ctypes.byref(info), None, ctypes.c_void_p(stream))
    assert rc == 0, lib.rmpc_last_error()
    torch.cuda.synchronize(dev)
    return out


def test_static_scratch_entry_point_matches_the_handle_path(built):
    import torch
    from robot_mpcs_b200.planner.solver import Solver
    pm = yaml.safe_load(open(os.path.join(SOLVERS, SDIR, "paramMap.yaml")))
    torch.cuda.set_device(0)
    solver = Solver.from_directory(os.path.join(SOLVERS, SDIR), device=0, max_batch=600)
    lib = solver.lib
    assert lib.rmpc_static_max_batch() == 16384
    # batch 1: 600 feasible starts.  batch 2: 200 problems, the odd ones warm-started from a plan an obstacle moved onto
    s1 = sc.boxer_radial(pm, 600, seed=31)
    d1 = {k: torch.as_tensor(v, device="cuda:0") for k, v in s1.items()}
    s2 = sc.boxer_radial(pm, 200, seed=32)
    first = solver.solve_batch(*(torch.as_tensor(s2[k], device="cuda:0") for k in ("xinit", "x0", "params")))
    z1 = first["z"].cpu().numpy()
    ee = z1[:, :, 0:2] + 0.4 * np.stack([np.cos(z1[:, :, 2]), np.sin(z1[:, :, 2])], 2)
    t = ee[:, 19] - ee[:, 17]
    t /= np.linalg.norm(t, axis=1, keepdims=True) + 1e-12
    c = ee[:, 18] + 0.85 * np.stack([-t[:, 1], t[:, 0]], 1)
    odd = np.arange(200) % 2 == 1
    oi = pm["obst"][0]
    P = s2["params"].copy()
    P[odd, :, oi:oi + 2] = c[odd, None, :]
    P[odd, :, oi + 2] = 0.0
    P[odd, :, oi + 3] = 0.3
    x0 = s2["x0"].copy()
    x0[odd] = z1[odd]
    d2 = {"xinit": torch.as_tensor(s2["xinit"], device="cuda:0"), "x0": torch.as_tensor(x0, device="cuda:0"), "params": torch.as_tensor(P, device="cuda:0")}
    for d in (d1, d2, d1):                                     # two calls in a row, and the first again
        B = d["xinit"].shape[0]
        ref = solver.solve_batch(d["xinit"], d["x0"], d["params"])
        torch.cuda.synchronize()
        got = _call_static(lib, d, B, solver.N, solver.nvar, solver.nu)
        ef = ref["exitflag"].cpu().numpy()
        assert np.array_equal(got["exitflag"].cpu().numpy(), ef) and (got["exitflag"] != -1).all()     # every problem written
        assert np.array_equal(got["iters"].cpu().numpy(), ref["iters"].cpu().numpy())
        assert (ef == 1).mean() > 0.85
        zr, zg = ref["z"].cpu().numpy(), got["z"].cpu().numpy()
        rel = np.abs(zg - zr).max(axis=(1, 2)) / np.abs(zr).max(axis=(1, 2))
        assert rel.max() < 1e-9, rel.max()                       # the same kernels on the same problems
        assert np.allclose(got["action"].cpu().numpy(), ref["action"].cpu().numpy(), rtol=0, atol=1e-9)
        ok = ef == 1
        assert np.allclose(got["obj"].cpu().numpy()[ok], ref["obj"].cpu().numpy()[ok], rtol=1e-9)
    # beyond the stated maximum: refused, nothing launched
    n0 = lib.rmpc_launch_count()
    rc = lib.rmpc_solve_batch_static(16385, d1["xinit"].data_ptr(), d1["x0"].data_ptr(), d1["params"].data_ptr(), None, None, None, None, None)
    assert rc != 0 and b"16384" in lib.rmpc_last_error() and lib.rmpc_launch_count() == n0
