"""ORACLE (test infrastructure only, never on the product path): numpy restatement of the reference's free-space
decomposition, robotmpcs/utils/free_space_decomposition.py, as driven by examples/boxer_example.py.

PINNED: tests/golden/fsd_reference.json holds outputs of the reference's own `FreeSpaceDecomposition` class, imported
from /root/reference by tests/golden/make_fsd_golden.py (with a stub for its unused `urdfenvs` import);
tests/test_fsd.py checks this restatement against them.
"""
import numpy as np


def lidar_position(state, lidar_offset=0.4, lidar_z=0.02):
    """examples/boxer_example.py:138-144 (compute_constraints): seed point of the decomposition."""
    angle = state[2]
    rot = np.array([[np.cos(angle), -np.sin(angle)], [np.sin(angle), np.cos(angle)]])
    p = np.dot(rot, np.array([lidar_offset, 0.0])) + np.asarray(state[0:2], dtype=float)
    return np.array([p[0], p[1], lidar_z])


def halfplane(point, position):
    """HalfPlane.__init__ / constraint(), free_space_decomposition.py:11-14,52-53: [normal, constant]."""
    normal = position - point
    return np.concatenate((normal, [-np.dot(normal, point)]))


def compute_constraints(points, position, number_constraints, max_radius):
    """FreeSpaceDecomposition.compute_constraints (:80-99) followed by asdict (:104-116) -> [K, 4]."""
    dists = np.linalg.norm(points - position, axis=1)
    idx = np.argsort(dists, kind="stable")
    pts = points[idx]
    pts = pts[dists[idx] < max_radius]
    rows = []
    while pts.size > 0 and len(rows) < number_constraints:
        c = halfplane(pts[0], position)
        rows.append(c)
        # point_infront_plane (:16-20), point by point with the same np.dot that made the constant: the chosen point
        # evaluates to exactly 0 and is dropped
        mask = np.array([np.dot(c[:3], q) + c[3] > 0 for q in pts])
        pts = pts[mask]
    n_real = len(rows)
    dummy = halfplane(position + np.array([20.0, 20.0, 0.0]), position)
    while len(rows) < number_constraints:
        rows.append(dummy)
    return np.array(rows), n_real


def batch(points, ref, number_constraints, max_radius, lidar_offset=0.4, lidar_z=0.02):
    """points[B,P,3], ref[B,N,>=3] -> [B,N,K,4]: the per-stage loop of boxer_example.py:191-201."""
    B, N = ref.shape[:2]
    out = np.zeros((B, N, number_constraints, 4))
    for b in range(B):
        for k in range(N):
            out[b, k], _ = compute_constraints(points[b], lidar_position(ref[b, k], lidar_offset, lidar_z), number_constraints, max_radius)
    return out
