"""Configuration schema, dimensions and the paramMap allocator of an MPC model.

Host-side mirror of robotmpcs/models/mpcBase.py:7-80: same dataclass fields, same dimension rules, same
first-come parameter allocation.  Instead of holding a symbolic FK object it holds the flattened
kinematic tree that the generator bakes into the CUDA specialisation.
"""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import Dict, List

from robot_mpcs_b200.utils.urdf import KinematicTree


@dataclass
class MpcConfiguration:  # mpcBase.py:7-22
    time_horizon: int
    time_step: float
    weights: dict
    slack: bool
    interval: int
    constraints: list
    objectives: list
    number_obstacles: int
    model_name: str
    initialization: str
    n: int
    control_mode: str
    name: str = "mpc"
    debug: bool = False


@dataclass
class RobotConfiguration:  # mpcBase.py:24-31
    collision_links: list
    selfCollision: dict
    urdf_file: str
    root_link: str
    end_link: str
    base_type: str


ASSET_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "assets")


def resolve_urdf(path: str) -> str:
    if os.path.isabs(path) and os.path.isfile(path):
        return path
    for cand in (path, os.path.join(ASSET_DIR, path), os.path.join(ASSET_DIR, os.path.basename(path))):
        if os.path.isfile(cand):
            return cand
    raise FileNotFoundError(path)


class MpcBase(object):
    def __init__(self, **kwargs):
        self._config = MpcConfiguration(**kwargs["mpc"])
        self._debug = kwargs.get("example", {}).get("debug", False)
        self._robot_config = RobotConfiguration(**kwargs["robot"])
        if self._config.slack:
            raise NotImplementedError("slack: True is dead code in the reference (ns is never set; SURVEY 3.6-6)")
        with open(resolve_urdf(self._robot_config.urdf_file), "r") as f:
            self._tree = KinematicTree(f.read())
        self._modelName = self._config.model_name
        n_chain = self._tree.n_movable(self._robot_config.root_link, self._robot_config.end_link)
        self._m = 3
        self._dt = self._config.time_step
        if self._robot_config.base_type == "holonomic":      # mpcBase.py:54-57
            self._n = n_chain
            self._nx = 2 * self._n
            self._nu = self._n
        elif self._robot_config.base_type == "diffdrive":    # mpcBase.py:58-61
            if n_chain != 0:
                raise NotImplementedError("diffdrive base with arm joints: the reference dynamics hard-code n=3 "
                                          "(diff_drive_mpc_model.py:24-31)")
            self._n = n_chain + 3
            self._nx = 2 * self._n + 2
            self._nu = 2 + n_chain
        else:
            raise ValueError(self._robot_config.base_type)
        self._ns = 0
        self._m_obst = 3
        self._N = self._config.time_horizon

    def addEntry2ParamMap(self, name: str, n_par: int):  # mpcBase.py:68-71
        if name not in self._paramMap:
            self._paramMap[name] = list(range(self._npar, self._npar + n_par))
            self._npar += n_par
