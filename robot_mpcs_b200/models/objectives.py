"""Objective plugins.  Mirrors robotmpcs/models/objectives/{ObjectiveManager,goal_reaching,constraint_avoidance}.py:
same names, same parameter allocation order (wu first, ObjectiveManager.py:14)."""
from __future__ import annotations


class GoalReaching:  # goal_reaching.py:10-34
    def __init__(self, model):
        self.model = model

    def set_parameters(self):
        m = self.model
        m.addEntry2ParamMap("goal", m._m)
        m.addEntry2ParamMap("wgoal", m._m)

    def describe(self, rb):
        m = self.model
        rb.goal(m._robot_config.end_link, m._paramMap["goal"][0], m._paramMap["wgoal"][0])


class ConstraintAvoidance:  # constraint_avoidance.py:14-32: N * w_i / (first row of module i)
    def __init__(self, model):
        self.model = model

    def set_parameters(self):
        m = self.model
        m.addEntry2ParamMap("wconstr", len(m._config.constraints))

    def describe(self, rb):
        m = self.model
        for i, first_row in enumerate(rb.module_first_row):
            if first_row is not None:
                rb.avoid(first_row, m._paramMap["wconstr"][i])


REGISTRY = {c.__name__: c for c in (GoalReaching, ConstraintAvoidance)}
