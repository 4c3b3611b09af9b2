"""Minimal URDF kinematic-tree reader used by the solver generator.

Replaces the part of forwardkinematics.GenericURDFFk (third party, absent) that the reference touches:
`GenericURDFFk(urdf, root, end, base_type)`, `.n()`, `.fk(q, root, link, positionOnly=True)`
(call sites robotmpcs/models/mpcBase.py:46-51,89-94).  Here the tree is not evaluated symbolically; it is
flattened into constant tables (rotation/translation/axis per joint) that the CUDA device functions in
csrc/rmpc_model.cuh consume as immediates.
"""
from __future__ import annotations

import math
import xml.etree.ElementTree as ET
from dataclasses import dataclass
from typing import Dict, List, Tuple


@dataclass
class Joint:
    name: str
    kind: int            # 0 fixed, 1 revolute/continuous, 2 prismatic
    parent_link: str
    child_link: str
    R: List[float]       # row-major 3x3 rotation of the joint origin
    t: List[float]
    axis: List[float]    # unit axis in the joint frame


def _rot_rpy(r: float, p: float, y: float) -> List[float]:
    cr, sr, cp, sp, cy, sy = math.cos(r), math.sin(r), math.cos(p), math.sin(p), math.cos(y), math.sin(y)
    return [cy * cp, cy * sp * sr - sy * cr, cy * sp * cr + sy * sr,
            sy * cp, sy * sp * sr + cy * cr, sy * sp * cr - cy * sr,
            -sp, cp * sr, cp * cr]


def _snap(v: float) -> float:
    """URDFs write pi/2 with ~1e-11 error; snap entries that are 0/±1 up to 1e-9 so they fold at compile time."""
    for c in (0.0, 1.0, -1.0):
        if abs(v - c) < 1e-9:
            return c
    return v


class KinematicTree:
    def __init__(self, urdf_text: str):
        robot = ET.fromstring(urdf_text)
        self.by_child: Dict[str, Joint] = {}
        kinds = {"fixed": 0, "revolute": 1, "continuous": 1, "prismatic": 2}
        for j in robot.findall("joint"):
            org = j.find("origin")
            xyz = [float(v) for v in (org.get("xyz", "0 0 0") if org is not None else "0 0 0").split()]
            rpy = [float(v) for v in (org.get("rpy", "0 0 0") if org is not None else "0 0 0").split()]
            ax = j.find("axis")
            axis = [float(v) for v in (ax.get("xyz") if ax is not None else "1 0 0").split()]
            nrm = math.sqrt(sum(a * a for a in axis)) or 1.0
            jt = j.get("type")
            if jt not in kinds:
                raise ValueError(f"unsupported joint type {jt}")
            self.by_child[j.find("child").get("link")] = Joint(
                j.get("name"), kinds[jt], j.find("parent").get("link"), j.find("child").get("link"),
                [_snap(v) for v in _rot_rpy(*rpy)], xyz, [a / nrm for a in axis])
        self.links = [l.get("name") for l in robot.findall("link")]

    def chain(self, root: str, link: str) -> List[Joint]:
        out, cur = [], link
        while cur != root:
            if cur not in self.by_child:
                raise KeyError(f"link '{link}' is not a descendant of '{root}'")
            jt = self.by_child[cur]
            out.append(jt)
            cur = jt.parent_link
        return out[::-1]

    def n_movable(self, root: str, end: str) -> int:
        return sum(1 for j in self.chain(root, end) if j.kind != 0)

    def positions(self, q, root: str, end: str, links):
        """Numeric forward kinematics (numpy, batched): q[B, n] over the movable joints of the root->end chain, in chain
        order -> {link: position[B, 3] in the root frame}.  The same chain product the generated device code evaluates
        (csrc/rmpc_model.cuh kin_frames); used by the scenario generators to reject states that start in collision."""
        import numpy as np
        q = np.asarray(q, dtype=np.float64)
        qi = {j.name: i for i, j in enumerate(jt for jt in self.chain(root, end) if jt.kind != 0)}
        out = {}
        for link in links:
            R = np.broadcast_to(np.eye(3), (q.shape[0], 3, 3)).copy()
            t = np.zeros((q.shape[0], 3))
            for jt in self.chain(root, link):
                t = t + R @ np.array(jt.t)
                R = R @ np.array(jt.R).reshape(3, 3)
                if jt.kind == 0:
                    continue
                a = np.array(jt.axis)
                v = q[:, qi[jt.name]]
                if jt.kind == 2:
                    t = t + (R @ a) * v[:, None]
                else:
                    K = np.array([[0, -a[2], a[1]], [a[2], 0, -a[0]], [-a[1], a[0], 0]])
                    Rj = np.eye(3)[None] + np.sin(v)[:, None, None] * K[None] + (1 - np.cos(v))[:, None, None] * (K @ K)[None]
                    R = R @ Rj
            out[link] = t
        return out
