"""Generates tests/golden/fsd_reference.json by running the REFERENCE's own FreeSpaceDecomposition class
(/root/reference/robotmpcs/utils/free_space_decomposition.py) on seeded point clouds.  Runs only in the build container
(the reference tree does not travel); the module's unused `urdfenvs` import is satisfied by an empty stub.

    python tests/golden/make_fsd_golden.py
"""
import importlib.util
import json
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from robot_mpcs_b200 import scenarios as sc  # noqa: E402

REF = "/root/reference/robotmpcs/utils/free_space_decomposition.py"


def load_reference_class():
    for name in ("urdfenvs", "urdfenvs.urdf_common", "urdfenvs.urdf_common.helpers"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.modules["urdfenvs.urdf_common.helpers"].add_shape = None
    sys.modules["urdfenvs.urdf_common.helpers"].quaternion_between_vectors = None
    spec = importlib.util.spec_from_file_location("ref_fsd", REF)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod.FreeSpaceDecomposition


def main():
    FSD = load_reference_class()
    cases = []
    for seed, (B, N, P, K, R) in enumerate([(3, 4, 64, 8, 5.0), (2, 3, 64, 5, 5.0), (2, 2, 17, 8, 1.5), (1, 2, 100, 16, 5.0), (1, 1, 64, 4, 0.05)]):
        pts, ref = sc.lidar_scene(B, N, P, seed=seed)
        out = np.zeros((B, N, K, 4))
        n_real = np.zeros((B, N), dtype=int)
        for b in range(B):
            for k in range(N):
                # examples/boxer_example.py:132-148
                angle = ref[b, k, 2]
                rot = np.array([[np.cos(angle), -np.sin(angle)], [np.sin(angle), np.cos(angle)]])
                pl = np.dot(rot, np.array([0.4, 0.0])) + ref[b, k, 0:2]
                fsd = FSD(number_constraints=K, max_radius=R)
                fsd.set_position(np.array([pl[0], pl[1], 0.02]))
                fsd.compute_constraints(pts[b])
                out[b, k] = np.array(list(fsd.asdict().values()))
                n_real[b, k] = len(fsd.constraints())
        cases.append({"seed": seed, "B": B, "N": N, "P": P, "K": K, "max_radius": R, "out": out.tolist(), "n_real": n_real.tolist()})
    with open(os.path.join(HERE, "fsd_reference.json"), "w") as f:
        json.dump({"source": REF, "generator": "robot_mpcs_b200.scenarios.lidar_scene", "cases": cases}, f)
    print("wrote", len(cases), "cases")


if __name__ == "__main__":
    main()
