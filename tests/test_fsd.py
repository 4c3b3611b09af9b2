"""Free-space decomposition.  Parity is PINNED here: tests/golden/fsd_reference.json was produced
by the reference's own FreeSpaceDecomposition class (tests/golden/make_fsd_golden.py).  CPU: the oracle restatement
against those vectors.  GPU: the CUDA kernel, through the C ABI, against the vectors and against the oracle on larger
seeded batches; plane selection must be identical, coefficients equal to rounding."""
import json
import os

import numpy as np
import pytest
import yaml

from conftest import GOLDEN, SOLVERS
from oracle import fsd as ofsd
from robot_mpcs_b200 import scenarios as sc

SDIR = "boxer_n3_01_H40_noSlack"   # the supermarket model: 8 half-planes per stage, N = 40


def _golden():
    return json.load(open(os.path.join(GOLDEN, "fsd_reference.json")))["cases"]


def test_oracle_matches_reference_vectors():
    for c in _golden():
        pts, ref = sc.lidar_scene(c["B"], c["N"], c["P"], seed=c["seed"])
        out = ofsd.batch(pts, ref, c["K"], c["max_radius"])
        assert np.array_equal(out, np.array(c["out"])), c["seed"]       # same numpy operations: same bits


def test_oracle_properties():
    """Every kept point lies behind at least one returned plane or outside the radius; the seed is strictly inside."""
    pts, ref = sc.lidar_scene(4, 3, 64, seed=7)
    out = ofsd.batch(pts, ref, 16, 5.0)
    for b in range(4):
        for k in range(3):
            p = ofsd.lidar_position(ref[b, k])
            assert (out[b, k, :, :3] @ p + out[b, k, :, 3] > 0).all()
            near = np.linalg.norm(pts[b] - p, axis=1) < 5.0
            vals = pts[b][near] @ out[b, k, :, :3].T + out[b, k, :, 3]
            assert (vals.min(axis=1) <= 1e-12).all()                     # 16 planes suffice for 64 rays here


def _solver(B):
    from robot_mpcs_b200.planner.solver import Solver
    return Solver.from_directory(os.path.join(SOLVERS, SDIR), device=0, max_batch=B)


@pytest.mark.gpu
def test_kernel_matches_reference_vectors(built):
    import torch
    solver = _solver(4)
    for c in _golden():
        pts, ref = sc.lidar_scene(c["B"], solver.N, c["P"], seed=c["seed"])   # same seed: the first c["N"] stages coincide
        pts_g, ref_g = sc.lidar_scene(c["B"], c["N"], c["P"], seed=c["seed"])
        ref[:, :c["N"]] = ref_g
        assert np.array_equal(pts, pts_g)
        out = solver.fsd_constraints(torch.as_tensor(pts, device="cuda:0"), torch.as_tensor(ref, device="cuda:0"), c["K"], c["max_radius"])
        got = out.cpu().numpy()[:, :c["N"]]
        want = np.array(c["out"])
        assert np.array_equal(np.abs(got[..., :3]) == 20.0, np.abs(want[..., :3]) == 20.0)   # same number of real planes
        assert np.abs(got - want).max() <= 1e-12 * max(1.0, np.abs(want).max()), c["seed"]


@pytest.mark.gpu
def test_kernel_vs_oracle_batch_and_param_scatter(built):
    """B = 64 problems x 40 stages x 8 planes: against the oracle; written straight into the parameter array at the
    paramMap indices; reference states taken from a plan z[B,N,nvar] by strides; one state per problem by stride 0."""
    import torch
    B, K = 64, 8
    solver = _solver(B)
    pm = yaml.safe_load(open(os.path.join(SOLVERS, SDIR, "paramMap.yaml")))
    idx = [pm[f"lin_constrs_{i}"][0] for i in range(K)]
    pts, ref = sc.lidar_scene(B, solver.N, 64, seed=21)
    want = ofsd.batch(pts, ref, K, 5.0)
    z = torch.zeros(B, solver.N, solver.nvar, dtype=torch.float64, device="cuda:0")
    z[:, :, :3] = torch.as_tensor(ref, device="cuda:0")
    params = torch.full((B, solver.N, solver.npar), -7.0, dtype=torch.float64, device="cuda:0")
    out = solver.fsd_constraints(torch.as_tensor(pts, device="cuda:0"), z, K, 5.0, params=params, idx=idx)
    got = out.cpu().numpy()
    assert np.abs(got - want).max() <= 1e-12 * np.abs(want).max()
    P = params.cpu().numpy()
    for c in range(K):
        assert np.array_equal(P[:, :, idx[c]:idx[c] + 4], got[:, :, c])
    untouched = np.ones(solver.npar, dtype=bool)
    for c in range(K):
        untouched[idx[c]:idx[c] + 4] = False
    assert (P[:, :, untouched] == -7.0).all()
    one = solver.fsd_constraints(torch.as_tensor(pts, device="cuda:0"), torch.as_tensor(ref[:, 0], device="cuda:0"), K, 5.0).cpu().numpy()
    assert np.array_equal(one, np.repeat(got[:, :1], solver.N, axis=1))


@pytest.mark.gpu
def test_reference_shaped_class(built):
    from robot_mpcs_b200.utils.free_space_decomposition import FreeSpaceDecomposition
    c = _golden()[0]
    pts, ref = sc.lidar_scene(c["B"], c["N"], c["P"], seed=c["seed"])
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        FreeSpaceDecomposition(number_constraints=c["K"], max_radius=c["max_radius"])
    f = FreeSpaceDecomposition(number_constraints=c["K"], max_radius=c["max_radius"], solver=_solver(1))
    f.set_position(ofsd.lidar_position(ref[1, 2]))
    f.compute_constraints(pts[1])
    want = np.array(c["out"])[1, 2]
    assert np.abs(np.array(list(f.asdict().values())) - want).max() < 1e-12 * np.abs(want).max()
    assert len(f.constraints()) == c["n_real"][1][2]
    assert f.aslist().shape == (c["K"], 4) and (f.aslist()[len(f.constraints()):, :2] == 20.0).all()
    assert all(abs(np.dot(h.normal(), h.point()) + h.constant()) < 1e-12 for h in f.constraints())   # planes pass through their points
