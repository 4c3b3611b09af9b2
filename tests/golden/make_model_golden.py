"""Golden vectors for the NLP definition, produced by the REFERENCE'S OWN model classes:
/root/reference/robotmpcs/models/{mpcBase,mpcModel,diff_drive_mpc_model}.py, inequalities/*, objectives/* and
utils/utils.py are imported and their `eval_inequalities(z, p)`, `eval_objectives(z, p)` and `continuous_dynamics(x, u)`
are called on seeded numeric points; their paramMap / npar / number_inequalities are recorded as built by their own
constructors.

What is stubbed, and why that does not weaken the pin:
  casadi             absent offline.  The reference uses ten of its functions (dot, mtimes, norm_2, vertcat, vcat, cos, sin,
                     fabs, SX(...), .size()) as plain arithmetic on symbols; here they are the same arithmetic on numbers.
  forwardkinematics  absent offline.  GenericURDFFk is replaced by oracle/nlp.py:ChainFk on the same URDF text, so link
                     positions are NOT pinned by this file (they are by the FK anchors of SURVEY.md §3.5,
                     tests/test_oracle_model.py); everything built on top of them is.
  forcespro          only constructs SymbolicModel / CodeOptions; not called here.
One line of the reference cannot run as written: RadialConstraints.eval_constraint passes an undefined name `j`
(inequalities/RadialConstraints.py:22, NameError).  It is called here as `eval_obstacleDistances(z, p)`, the two-argument
method it wraps (mpcBase.py:82); the generator records that the original raises.

Run in the build container only (needs /root/reference):  python tests/golden/make_model_golden.py
-> tests/golden/model_reference.json
"""
import importlib.metadata
import json
import os
import sys
import types

import numpy as np
import yaml

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
CONFIG = os.path.join(ROOT, "robot_mpcs_b200", "config")
ASSETS = os.path.join(ROOT, "robot_mpcs_b200", "assets")
CASES = ["pointRobotMpc.yaml", "boxerRadialMpc.yaml", "pandaMpc.yaml", "boxerSupermarketMpc.yaml", "boxerMpc.yaml",
         "boxerVelMpc.yaml"]   # the last one instantiates VelLimitConstraints (no shipped config of the reference does)


class V(np.ndarray):
    """Numbers shaped like CasADi SX: everything is a 2-D matrix, a vector is a column, an element is 1 x 1."""
    def __new__(cls, a):
        a = np.asarray(a, dtype=float)
        if a.ndim == 0:
            a = a.reshape(1, 1)
        elif a.ndim == 1:
            a = a.reshape(-1, 1)
        return a.view(cls)

    def __getitem__(self, key):
        if isinstance(key, tuple):
            return V(np.asarray(self)[key])
        return V(np.asarray(self).reshape(-1)[key])     # column vector: v[i], v[a:b], v[[i, j, ...]]

    def __setitem__(self, key, value):
        if isinstance(value, np.ndarray) and int(np.prod(value.shape)) == 1:
            value = float(np.asarray(value).reshape(-1)[0])
        np.ndarray.__setitem__(self, key, value)

    def size(self):  # casadi: (rows, columns)
        return tuple(self.shape)


def casadi_stub():
    ca = types.ModuleType("casadi")

    class SX:
        def __new__(cls, *a):
            if len(a) == 2:
                return V(np.zeros((int(a[0]), int(a[1]))))
            return V(np.atleast_1d(np.asarray(a[0], dtype=float)))
    ca.SX = SX
    flat = lambda x: np.asarray(x, dtype=float).reshape(-1)
    ca.dot = lambda a, b: V([float(flat(a) @ flat(b))])
    ca.mtimes = lambda A, b: V(np.asarray(A, dtype=float) @ flat(b))
    ca.norm_2 = lambda v: V([float(np.sqrt((flat(v) ** 2).sum()))])
    ca.vertcat = lambda *xs: V(np.concatenate([flat(x) for x in xs]))
    ca.vcat = lambda xs: V(np.concatenate([flat(x) for x in xs]))
    ca.cos = lambda x: V(np.cos(np.asarray(x, dtype=float)))
    ca.sin = lambda x: V(np.sin(np.asarray(x, dtype=float)))
    ca.fabs = lambda x: V(np.abs(np.asarray(x, dtype=float)))
    return ca


def install_stubs():
    import torch
    from oracle.nlp import ChainFk
    importlib.metadata.version = lambda name: "0.0.0"     # robotmpcs/__init__.py asks for an installed distribution
    sys.modules["casadi"] = casadi_stub()
    fp = types.ModuleType("forcespro")
    fp.nlp = types.ModuleType("forcespro.nlp")
    fp.nlp.SymbolicModel = lambda N: types.SimpleNamespace(N=N)
    fp.CodeOptions = lambda name: types.SimpleNamespace(nlp=types.SimpleNamespace(integrator=types.SimpleNamespace()))
    sys.modules["forcespro"], sys.modules["forcespro.nlp"] = fp, fp.nlp

    class GenericURDFFk:
        def __init__(self, urdf, root_link, end_link, base_type="holonomic"):
            self._c = ChainFk(urdf, root_link, end_link, base_type)

        def n(self):
            return self._c.n()

        def fk(self, q, root, link, positionOnly=True):
            return V(self._c.fk(torch.as_tensor(np.asarray(q, dtype=float).reshape(-1)), root, link).numpy())
    for name in ("forwardkinematics", "forwardkinematics.fksCommon", "forwardkinematics.fksCommon.fk",
                 "forwardkinematics.urdfFks", "forwardkinematics.urdfFks.generic_urdf_fk"):
        sys.modules[name] = types.ModuleType(name)
    sys.modules["forwardkinematics.fksCommon.fk"].ForwardKinematics = object
    sys.modules["forwardkinematics.urdfFks.generic_urdf_fk"].GenericURDFFk = GenericURDFFk
    for name in ("pybullet", "urdfenvs", "urdfenvs.sensors", "urdfenvs.sensors.free_space_decomposition"):   # utils/utils.py plotting helpers
        sys.modules.setdefault(name, types.ModuleType(name))


def sample(rng, model, paramMap, npar):
    """A numeric (z, p) away from every singularity (positive radii, non-zero plane normals, weights)."""
    nvar = model._nx + model._nu
    z = rng.uniform(-1.0, 1.0, nvar)
    p = rng.uniform(-2.0, 2.0, npar)
    for name, idx in paramMap.items():
        if name in ("wgoal", "wu", "wconstr", "r_body"):
            p[idx] = rng.uniform(0.1, 1.0, len(idx))
        if name == "obst":
            p[idx[3::4]] = rng.uniform(0.1, 0.5, len(idx) // 4)
    return z, p


def main():
    install_stubs()
    sys.path.insert(0, "/root/reference")
    from robotmpcs.models.diff_drive_mpc_model import MpcDiffDriveModel
    from robotmpcs.models.mpcModel import MpcModel
    from robotmpcs.models.inequalities.RadialConstraints import RadialConstraints
    radial_raises = None
    golden = {}
    os.chdir(ASSETS)   # the configs name the URDF by a bare file name (examples/makeSolver.py:17 prefixes the assets directory)
    for cfg in CASES:
        setup = yaml.safe_load(open(os.path.join(CONFIG, cfg)))
        cls = MpcModel if setup["robot"]["base_type"] == "holonomic" else MpcDiffDriveModel   # examples/makeSolver.py:18-21
        model = cls(initParamMap=True, **setup)
        model.setModel()
        model.setCodeoptions()
        rng = np.random.default_rng(11)
        if "RadialConstraints" in setup["mpc"]["constraints"] and radial_raises is None:
            z, p = sample(rng, model, model._paramMap, model._npar)
            try:
                model._inequality_manager.eval_inequalities(V(z), V(p))
                radial_raises = "ran"
            except NameError as e:
                radial_raises = "NameError: " + str(e)
        RadialConstraints.eval_constraint = lambda self, z, p: self.eval_obstacleDistances(z, p)   # see the module docstring
        samples = []
        for _ in range(6):
            z, p = sample(rng, model, model._paramMap, model._npar)
            h = model._model.ineq(V(z), V(p))
            J = model._model.objective(V(z), V(p))
            JN = model._model.objectiveN(V(z), V(p))
            xd = model._model.continuous_dynamics(V(z[:model._nx]), V(z[model._nx:]))
            samples.append({"z": z.tolist(), "p": p.tolist(), "h": [float(np.asarray(r).reshape(-1)[0]) for r in h],
                            "J": float(np.asarray(J).reshape(-1)[0]), "JN": float(np.asarray(JN).reshape(-1)[0]),
                            "xdot": np.asarray(xd, dtype=float).reshape(-1).tolist()})
        golden[cfg] = {"paramMap": model._paramMap, "npar": model._npar, "nh": model.number_inequalities, "model_nh": int(model._model.nh),
                       "nx": model._nx, "nu": model._nu, "n": model._n, "nvar": int(model._model.nvar), "neq": int(model._model.neq),
                       "lb": np.asarray(model._model.lb).tolist(), "ub": np.asarray(model._model.ub).tolist(),
                       "E": np.asarray(model._model.E).tolist(), "solverName": model._solverName,
                       "integrator": [model._codeoptions.nlp.integrator.type, model._codeoptions.nlp.integrator.Ts, model._codeoptions.nlp.integrator.nodes],
                       "samples": samples}
        print(cfg, "npar", model._npar, "nh", model.number_inequalities, "rows returned", len(samples[0]["h"]), "J", samples[0]["J"])
    golden["RadialConstraints.eval_constraint as written"] = radial_raises
    print("RadialConstraints.eval_constraint as written:", radial_raises)
    with open(os.path.join(ROOT, "tests", "golden", "model_reference.json"), "w") as f:
        json.dump(golden, f)


if __name__ == "__main__":
    main()
