"""Free-space decomposition with the reference's interface (robotmpcs/utils/free_space_decomposition.py:61-139), computed
by the CUDA kernel of csrc/rmpc_fsd.cuh through the C ABI (`rmpc_fsd_constraints`).  There is no CPU fallback: the class
needs a solver handle (any generated solver library exports the entry point).

    reference                                           here
    FreeSpaceDecomposition(number_constraints, max_radius)    FreeSpaceDecomposition(number_constraints, max_radius, solver=...)
    .set_position(p); .compute_constraints(points)            same (one seed point: B = 1, one stage)
    .constraints() / .asdict() / .aslist()                    same shapes
    —                                                         compute_constraints_batch(points[B,P,3], ref[B,N,3]) -> [B,N,K,4]
"""
from __future__ import annotations

import numpy as np

try:
    import torch
except Exception:  # pragma: no cover
    torch = None


class HalfPlane:
    """Value holder with the reference's accessors (free_space_decomposition.py:6-53)."""

    def __init__(self, coefficients: np.ndarray, point=None):
        self._normal_vector = np.asarray(coefficients[:3], dtype=float)
        self._constant = float(coefficients[3])
        self._point = point

    def normal(self):
        return self._normal_vector

    def constant(self):
        return self._constant

    def point(self):
        return self._point

    def point_behind_plane(self, point):
        return float(np.dot(self.normal(), point) + self.constant()) <= 0

    def point_infront_plane(self, point):
        return not self.point_behind_plane(point)

    def constraint(self):
        return np.concatenate((self.normal(), np.array([self.constant()])))


class FreeSpaceDecomposition:
    def __init__(self, number_constraints: int = 10, max_radius: float = 1.0, solver=None, device: int = 0):
        if solver is None:
            raise RuntimeError("FreeSpaceDecomposition runs on the GPU: pass solver=<robot_mpcs_b200.planner.solver.Solver> "
                               "(or planner._solver); there is no CPU fallback")
        self._number_constraints = number_constraints
        self._max_radius = max_radius
        self._solver = solver
        self._device = torch.device("cuda", device)
        self._constraints = []
        self._rows = None
        self._position = None

    def set_position(self, position: np.ndarray):
        self._position = np.asarray(position, dtype=float)

    def compute_constraints(self, points: np.ndarray):
        """One seed point (set_position) against one cloud.  The kernel derives the seed from a state (x, y, theta) as
        (x, y) + Rot(theta) (offset, 0); with offset 0 the state is the seed itself."""
        p = self._position
        pts = torch.as_tensor(np.ascontiguousarray(points, dtype=np.float64)[None], device=self._device)
        ref = torch.as_tensor(np.array([[p[0], p[1], 0.0]]), device=self._device)
        out = self._solver.fsd_constraints(pts, ref, self._number_constraints, self._max_radius, lidar_offset=0.0,
                                           lidar_z=float(p[2]) if p.shape[0] > 2 else 0.0)
        self._rows = out[0, 0].cpu().numpy()
        dummy_normal = np.array([-20.0, -20.0, 0.0])
        self._constraints = [HalfPlane(r, point=p - r[:3]) for r in self._rows if not np.array_equal(r[:3], dummy_normal)]

    def compute_constraints_batch(self, points, ref, params=None, idx=None, lidar_offset=0.4, lidar_z=0.02):
        """points[B,P,3], ref[B,N,>=3] or [B,>=3] (CUDA tensors) -> planes[B,N,K,4]; optionally scattered into params."""
        return self._solver.fsd_constraints(points, ref, self._number_constraints, self._max_radius, params=params, idx=idx,
                                            lidar_offset=lidar_offset, lidar_z=lidar_z)

    def constraints(self):
        return self._constraints

    def asdict(self):
        return {f"constraint_{i}": self._rows[i].copy() for i in range(self._number_constraints)}

    def aslist(self):
        """The reference's aslist() builds its dummy rows with the arguments of HalfPlane swapped (:126-128): normal
        (+20, +20, 0) through the seed point.  Kept."""
        out = self._rows.copy()
        for i in range(len(self._constraints), self._number_constraints):
            n = np.array([20.0, 20.0, 0.0])
            out[i] = np.concatenate((n, [-np.dot(n, self._position)]))
        return out
