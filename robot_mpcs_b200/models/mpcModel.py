"""MpcModel: config -> NLP description -> generated CUDA solver directory.

Host-side mirror of robotmpcs/models/mpcModel.py:14-141.  The reference hands CasADi callbacks to
forcespro.nlp.SymbolicModel and asks the FORCESPRO server for C code; here `setModel()` flattens the same
NLP (rows in config order, parameters in paramMap order, dynamics, bounds) into constant tables, and
`generateSolver(location)` emits `model_traits.h`, compiles csrc/rmpc_solver.cu against it with nvcc for
sm_100a, and writes `<location>/<solverName>/{lib<solverName>.so, paramMap.yaml, properties.yaml}` — the same
directory name and the same two yaml files the reference's planner loads (mpcPlanner.py:43-70).
"""
from __future__ import annotations

import json
import os
import shutil
import subprocess
from typing import Dict, List, Optional

import yaml

from robot_mpcs_b200.models.mpcBase import MpcBase
from robot_mpcs_b200.models import inequalities as _ineq
from robot_mpcs_b200.models import objectives as _obj

CSRC_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "csrc")
INCLUDE_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "include")
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "--expt-relaxed-constexpr", "-Xcompiler", "-fPIC", "-shared"]


class RowBuilder:
    """Collects the rows/objective terms the plugins declare and the kinematic sub-tree they need."""

    def __init__(self, model: "MpcModel"):
        self.m = model
        self.joints: List[dict] = []          # flattened tree joints (topological)
        self._joint_idx: Dict[str, int] = {}
        self.points: List[dict] = []          # {link, joint, mask}
        self._point_idx: Dict[str, int] = {}
        self.geo: List[dict] = []
        self.lim: List[dict] = []
        self.nh = 0
        self.goal_term: Optional[dict] = None
        self.ca: List[dict] = []
        self.module_first_row: List[Optional[int]] = []
        tree, rc = model._tree, model._robot_config
        self._q_of: Dict[str, int] = {}
        off = 0
        self._base_parent = -1
        if rc.base_type == "diffdrive":
            # planar base pose T(X, Y, theta) = prismatic x, prismatic y, revolute z in front of the chain
            I = [1.0, 0, 0, 0, 1.0, 0, 0, 0, 1.0]
            for nm, kind, axis, qi in (("__base_x", 2, [1.0, 0, 0], 0), ("__base_y", 2, [0, 1.0, 0], 1),
                                       ("__base_theta", 1, [0, 0, 1.0], 2)):
                self.joints.append({"name": nm, "parent": len(self.joints) - 1, "type": kind, "q": qi,
                                    "R": I, "t": [0.0, 0.0, 0.0], "axis": axis})
                self._joint_idx[nm] = len(self.joints) - 1
            self._base_parent = len(self.joints) - 1
            off = 3
        k = 0
        for jt in tree.chain(rc.root_link, rc.end_link):
            if jt.kind != 0:
                self._q_of[jt.name] = off + k
                k += 1

    # -- kinematics bookkeeping
    def _point(self, link: str) -> int:
        if link in self._point_idx:
            return self._point_idx[link]
        rc = self.m._robot_config
        chain = self.m._tree.chain(rc.root_link, link)
        parent = self._base_parent
        for jt in chain:
            if jt.name not in self._joint_idx:
                if jt.kind != 0 and jt.name not in self._q_of:
                    raise ValueError(f"link {link} depends on joint {jt.name} which is not on the root->end chain")
                self.joints.append({"name": jt.name, "parent": parent, "type": jt.kind,
                                    "q": self._q_of.get(jt.name, -1), "R": jt.R, "t": jt.t, "axis": jt.axis})
                self._joint_idx[jt.name] = len(self.joints) - 1
            parent = self._joint_idx[jt.name]
        mask, j = 0, parent
        while j >= 0:
            mask |= 1 << j
            j = self.joints[j]["parent"]
        self.points.append({"link": link, "joint": parent, "mask": mask})
        self._point_idx[link] = len(self.points) - 1
        return self._point_idx[link]

    # -- rows (called by the inequality plugins, in h order)
    def sphere(self, link, p_obst, p_rb):
        self.geo.append({"type": 0, "a": self._point(link), "b": -1, "p": p_obst, "rb": p_rb, "row": self.nh}); self.nh += 1

    def plane(self, link, p_plane, p_rb):
        self.geo.append({"type": 1, "a": self._point(link), "b": -1, "p": p_plane, "rb": p_rb, "row": self.nh}); self.nh += 1

    def pair(self, la, lb, p_rb):
        self.geo.append({"type": 2, "a": self._point(la), "b": self._point(lb), "p": -1, "rb": p_rb, "row": self.nh}); self.nh += 1

    def lower(self, var, p):
        self.lim.append({"var": var, "sign": 1, "p": p, "row": self.nh}); self.nh += 1

    def upper(self, var, p):
        self.lim.append({"var": var, "sign": -1, "p": p, "row": self.nh}); self.nh += 1

    # -- objective terms
    def goal(self, link, p_goal, p_wgoal):
        self.goal_term = {"pt": self._point(link), "p_goal": p_goal, "p_wgoal": p_wgoal}

    def avoid(self, row, p_w):
        geo = next((i for i, g in enumerate(self.geo) if g["row"] == row), -1)
        lim = next((i for i, l in enumerate(self.lim) if l["row"] == row), -1)
        self.ca.append({"row": row, "w": p_w, "geo": geo, "lim": lim})


def _arr(vals, fmt="{}"):
    vals = list(vals)
    if not vals:
        vals = [0]
    return "{" + ", ".join(fmt.format(v) for v in vals) + "}"


def _getter(name, typ, vals, fmt="{}"):
    return (f"  static constexpr RMPC_HD {typ} {name}(int i) {{ constexpr {typ} a[] = {_arr(vals, fmt)}; return a[i]; }}\n")


class MpcModel(MpcBase):
    def __init__(self, initParamMap=True, **kwargs):
        super().__init__(**kwargs)
        self._kwargs = kwargs
        if initParamMap:
            self.initParamMap()
        # constraints first, in config order (mpcModel.py:29-30) ...
        self.inequality_modules = [_ineq.REGISTRY[name](self) for name in self._config.constraints]
        for mod in self.inequality_modules:
            mod.set_parameters()
        # ... then wu and the objectives (ObjectiveManager.py:14, mpcModel.py:35-36)
        self.addEntry2ParamMap("wu", self._nu)
        self.objective_modules = [_obj.REGISTRY[name](self) for name in self._config.objectives]
        for mod in self.objective_modules:
            mod.set_parameters()
        self._rows: Optional[RowBuilder] = None
        self._solverName: Optional[str] = None
        self.number_inequalities = 0

    def initParamMap(self):  # mpcModel.py:40-42
        self._paramMap = {}
        self._npar = 0

    # ------------------------------------------------------------------------------------------
    def setModel(self):
        """Flatten the NLP (mpcModel.py:74-108): rows in module order, hl=0/hu=inf, E=[I 0], nvar, neq, xinitidx."""
        rb = RowBuilder(self)
        for mod in self.inequality_modules:
            first = rb.nh
            mod.rows(rb)
            rb.module_first_row.append(first if rb.nh > first else None)
        for mod in self.objective_modules:
            mod.describe(rb)
        if len(rb.joints) > 32:
            raise NotImplementedError("more than 32 joints on the used sub-tree")
        self._rows = rb
        self.number_inequalities = rb.nh
        return self

    def solverName(self) -> str:  # mpcModel.py:111-113
        name = self._modelName + "_n" + str(self._n) + "_" + str(self._dt).replace(".", "") + "_H" + str(self._N)
        if not self._config.slack:
            name += "_noSlack"
        return name

    def setCodeoptions(self, **kwargs):
        """mpcModel.py:110-126.  Integrator is fixed to ERK2 / Ts=dt / 5 nodes as there; `solverName=` overrides
        the directory name (the evident intent of mpcModel.py:114-115)."""
        self._solverName = kwargs.get("solverName", self.solverName())
        self._nvcc_extra = list(kwargs.get("nvcc_flags", []))
        return self

    # ------------------------------------------------------------------------------------------
    def traits_source(self) -> str:
        rb = self._rows
        assert rb is not None, "call setModel() first"
        name = self._solverName or self.solverName()
        f17 = "{!r}"
        src = ["// generated by robot_mpcs_b200.models.mpcModel — do not edit\n", "#pragma once\n",
               '#include "rmpc_model.cuh"\n', "struct ModelTraits {\n",
               f'  static constexpr const char* NAME = "{name}";\n',
               f"  static constexpr int N = {self._N}, NQ = {self._n}, NX = {self._nx}, NU = {self._nu}, "
               f"NVAR = {self._nx + self._nu}, NPAR = {self._npar}, NH = {rb.nh};\n",
               f"  static constexpr int BASE = {0 if self._robot_config.base_type == 'holonomic' else 1};\n",
               f"  static constexpr double DT = {self._dt!r};\n",
               f"  static constexpr int NJ = {len(rb.joints)}, NPT = {len(rb.points)}, NGEO = {len(rb.geo)}, "
               f"NLIM = {len(rb.lim)}, NCA = {len(rb.ca)};\n"]
        J = rb.joints
        src.append(_getter("j_parent", "int", [j["parent"] for j in J]))
        src.append(_getter("j_type", "int", [j["type"] for j in J]))
        src.append(_getter("j_q", "int", [j["q"] for j in J]))
        src.append("  static constexpr RMPC_HD double j_R(int j, int i) { constexpr double a[] = "
                   + _arr([v for j in J for v in j["R"]], f17) + "; return a[9 * j + i]; }\n")
        src.append("  static constexpr RMPC_HD double j_T(int j, int i) { constexpr double a[] = "
                   + _arr([v for j in J for v in j["t"]], f17) + "; return a[3 * j + i]; }\n")
        src.append("  static constexpr RMPC_HD double j_axis(int j, int i) { constexpr double a[] = "
                   + _arr([v for j in J for v in j["axis"]], f17) + "; return a[3 * j + i]; }\n")
        src.append(_getter("pt_joint", "int", [p["joint"] for p in rb.points]))
        src.append(_getter("pt_mask", "unsigned", [p["mask"] for p in rb.points], "{}u"))
        for key in ("type", "a", "b", "p", "rb", "row"):
            src.append(_getter("geo_" + key, "int", [g[key] for g in rb.geo]))
        for key in ("var", "sign", "p", "row"):
            src.append(_getter("lim_" + key, "int", [l[key] for l in rb.lim]))
        for key in ("row", "w", "geo", "lim"):
            src.append(_getter("ca_" + key, "int", [c[key] for c in rb.ca]))
        g = rb.goal_term
        src.append(f"  static constexpr int GOAL_PT = {g['pt'] if g else -1}, P_GOAL = {g['p_goal'] if g else 0}, "
                   f"P_WGOAL = {g['p_wgoal'] if g else 0}, P_WU = {self._paramMap['wu'][0]};\n")
        src.append("};\n")
        return "".join(src)

    def properties(self) -> dict:  # mpcModel.py:134
        return {"nx": self._nx, "nu": self._nu, "npar": self._npar, "ns": self._ns, "m": self._m,
                "constraints": list(self._config.constraints)}

    def generateSolver(self, location="./", compile_cuda=True, verbose=False):
        """Emit + compile the CUDA specialisation and write the solver directory (mpcModel.py:128-141)."""
        if self._rows is None:
            self.setModel()
        if self._solverName is None:
            self.setCodeoptions()
        if self._debug:
            location = os.path.join(location, "debug")
        out_dir = os.path.join(location, self._solverName)
        if os.path.isdir(out_dir):
            shutil.rmtree(out_dir)
        os.makedirs(out_dir)
        with open(os.path.join(out_dir, "model_traits.h"), "w") as f:
            f.write(self.traits_source())
        with open(os.path.join(out_dir, "paramMap.yaml"), "w") as f:
            yaml.dump(self._paramMap, f, default_flow_style=False)
        with open(os.path.join(out_dir, "properties.yaml"), "w") as f:
            yaml.dump(self.properties(), f, default_flow_style=False)
        with open(os.path.join(out_dir, "model.json"), "w") as f:
            json.dump({"name": self._solverName, "N": self._N, "n": self._n, "nx": self._nx, "nu": self._nu,
                       "npar": self._npar, "nh": self._rows.nh, "dt": self._dt,
                       "base_type": self._robot_config.base_type, "joints": self._rows.joints,
                       "points": self._rows.points, "geo": self._rows.geo, "lim": self._rows.lim,
                       "ca": self._rows.ca, "goal": self._rows.goal_term}, f, indent=1)
        if compile_cuda:
            compile_solver(out_dir, self._solverName, extra=getattr(self, "_nvcc_extra", []), verbose=verbose)
        return out_dir


def compile_solver(out_dir: str, name: str, extra=(), verbose=False) -> str:
    """nvcc -> <out_dir>/lib<name>.so for sm_100a (cross-compiles without a GPU)."""
    so = os.path.join(out_dir, f"lib{name}.so")
    cmd = ["nvcc", *NVCC_FLAGS, *extra, "-I", out_dir, "-I", CSRC_DIR, "-I", INCLUDE_DIR,
           os.path.join(CSRC_DIR, "rmpc_solver.cu"), "-o", so]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + res.stdout + res.stderr)
    if verbose:
        print(res.stdout + res.stderr)
    return so
