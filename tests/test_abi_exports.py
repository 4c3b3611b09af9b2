"""The C-ABI library of every generated solver directory loads on a CPU-only box and exports every symbol that
include/rmpc.h declares (no compute calls here: there is no CPU fallback to call)."""
import ctypes
import os
import re

import pytest

from conftest import ROOT, SOLVERS


def _declared():
    hdr = open(os.path.join(ROOT, "include", "rmpc.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    return sorted(set(re.findall(r"\b(rmpc_[a-z0-9_]+)\s*\(", hdr)))


def test_header_declares_the_abi():
    names = _declared()
    for n in ("rmpc_create", "rmpc_destroy", "rmpc_solve_batch", "rmpc_solve_batch_host", "rmpc_shift_horizon",
              "rmpc_set_params", "rmpc_last_error", "rmpc_get_dims", "rmpc_launch_count"):
        assert n in names


def test_every_solver_exports_every_symbol(built):
    from robot_mpcs_b200.planner.solver import EXPORTS, RmpcDims, load_library
    dirs = [d for d in sorted(os.listdir(SOLVERS)) if os.path.isdir(os.path.join(SOLVERS, d))]
    assert len(dirs) >= 5
    for d in dirs:
        lib = load_library(os.path.join(SOLVERS, d))
        for sym in _declared():
            assert hasattr(lib, sym), (d, sym)
        assert set(EXPORTS) == set(_declared())
        dims = RmpcDims()
        assert lib.rmpc_get_dims(ctypes.byref(dims)) == 0 and dims.nvar == dims.nx + dims.nu
        assert lib.rmpc_model_name().decode() == d


def test_no_gpu_fails_loudly(built):
    import torch
    if torch.cuda.is_available():
        pytest.skip("CUDA device present")
    from robot_mpcs_b200.planner.solver import Solver, SolverError
    with pytest.raises(SolverError, match="no CPU fallback|no usable CUDA"):
        Solver.from_directory(os.path.join(SOLVERS, "pointRobot_n3_005_H20_noSlack"), device=0, max_batch=1)


def test_exchange_region_layout_is_host_arithmetic(built):
    """rmpc_exchange_region_bytes needs no device: status | flags[depth][world] | acks[world] | counters[2][depth] | rows (256-byte
    aligned) of world * depth * rows_per_rank * (nu + 1) doubles (csrc/rmpc_exchange.cuh)."""
    from robot_mpcs_b200.planner.solver import RmpcDims, load_library
    lib = load_library(os.path.join(SOLVERS, "boxer_n3_01_H30_noSlack"))
    dims = RmpcDims()
    lib.rmpc_get_dims(ctypes.byref(dims))
    f = lib.rmpc_exchange_region_bytes
    for world, rows, depth in [(1, 1, 1), (2, 4096, 8), (8, 131072, 8), (16, 7, 3)]:
        head = 32 + 8 * depth * world + 8 * world + 8 * depth
        want = (head + 255) // 256 * 256 + 8 * depth * world * rows * (dims.nu + 1)
        assert f(world, rows, depth) == want
    assert f(0, 1, 1) == 0 and f(17, 1, 1) == 0 and f(2, 0, 1) == 0 and f(2, 1, 0) == 0
