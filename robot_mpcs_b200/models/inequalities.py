"""Inequality plugins h(z,p) >= 0.  Mirrors robotmpcs/models/inequalities/*.py: the same class names are
looked up from the `constraints:` list of the config (InequalityManager.py:15-23), each allocates the same
parameters in the same order, and each declares its rows in the same order.  Instead of building CasADi
expressions a plugin *describes* its rows to the generator (`RowBuilder`), which emits constant tables for
the fused CUDA evaluation in csrc/rmpc_model.cuh.
"""
from __future__ import annotations


class _Plugin:
    def __init__(self, model):
        self.model = model

    def set_parameters(self):
        raise NotImplementedError

    def rows(self, rb):
        raise NotImplementedError


class RadialConstraints(_Plugin):  # RadialConstraints.py:6-23 + mpcBase.py:82-101
    def set_parameters(self):
        m = self.model
        m.addEntry2ParamMap("r_body", 1)
        m.addEntry2ParamMap("obst", 4 * m._config.number_obstacles)

    def rows(self, rb):
        m = self.model
        for link in m._robot_config.collision_links:
            for i in range(m._config.number_obstacles):
                rb.sphere(link, m._paramMap["obst"][4 * i], m._paramMap["r_body"][0])


class LinearConstraints(_Plugin):  # LinearConstraints.py:8-40
    def set_parameters(self):
        m = self.model
        m.addEntry2ParamMap("r_body", 1)
        for i in range(m._config.number_obstacles):
            m.addEntry2ParamMap("lin_constrs_" + str(i), 4)

    def rows(self, rb):
        m = self.model
        for link in m._robot_config.collision_links:
            for i in range(m._config.number_obstacles):
                rb.plane(link, m._paramMap["lin_constrs_" + str(i)][0], m._paramMap["r_body"][0])


class SelfCollisionAvoidanceConstraints(_Plugin):  # SelfCollisionAvoidanceConstraints.py:5-27
    def set_parameters(self):
        self.model.addEntry2ParamMap("r_body", 1)

    def rows(self, rb):
        m = self.model
        for a, b in m._robot_config.selfCollision["pairs"]:
            rb.pair(a, b, m._paramMap["r_body"][0])


class JointLimitConstraints(_Plugin):  # JointLimitConstraints.py:5-31
    def set_parameters(self):
        m = self.model
        m.addEntry2ParamMap("lower_limits", m._n)
        m.addEntry2ParamMap("upper_limits", m._n)

    def rows(self, rb):
        m = self.model
        for j in range(m._n):
            rb.lower(j, m._paramMap["lower_limits"][j])
            rb.upper(j, m._paramMap["upper_limits"][j])


class InputLimitConstraints(_Plugin):  # InputLimitConstraints.py:4-29
    def set_parameters(self):
        m = self.model
        m.addEntry2ParamMap("lower_limits_u", m._nu)
        m.addEntry2ParamMap("upper_limits_u", m._nu)

    def rows(self, rb):
        m = self.model
        for j in range(m._nu):
            rb.lower(m._nx + j, m._paramMap["lower_limits_u"][j])
            rb.upper(m._nx + j, m._paramMap["upper_limits_u"][j])


class VelLimitConstraints(_Plugin):  # VelLimitConstraints.py:5-31 (4 rows: last two entries of qdot = z[n:nx])
    def set_parameters(self):
        m = self.model
        m.addEntry2ParamMap("lower_limits_vel", 2)
        m.addEntry2ParamMap("upper_limits_vel", 2)

    def rows(self, rb):
        m = self.model
        for j in range(2):
            rb.lower(m._nx - 2 + j, m._paramMap["lower_limits_vel"][j])
            rb.upper(m._nx - 2 + j, m._paramMap["upper_limits_vel"][j])


REGISTRY = {c.__name__: c for c in (RadialConstraints, LinearConstraints, SelfCollisionAvoidanceConstraints,
                                    JointLimitConstraints, InputLimitConstraints, VelLimitConstraints)}
