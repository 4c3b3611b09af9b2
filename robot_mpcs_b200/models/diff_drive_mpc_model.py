"""Differential-drive variant (boxer).  Mirror of robotmpcs/models/diff_drive_mpc_model.py:7-41: state
x = [X, Y, theta, Xdot, Ydot, thetadot, v, omega], input u = [v_dot, omega_dot], continuous dynamics
[cos(theta) v, sin(theta) v, omega, 0, 0, 0, u].  Dimensions come from MpcBase (base_type 'diffdrive');
the dynamics themselves live in csrc/rmpc_model.cuh (dyn_value / dyn_derivatives, BASE == 1)."""
from robot_mpcs_b200.models.mpcModel import MpcModel


class MpcDiffDriveModel(MpcModel):
    def __init__(self, initParamMap=True, **kwargs):
        if kwargs["robot"]["base_type"] != "diffdrive":
            raise ValueError("MpcDiffDriveModel needs robot.base_type == 'diffdrive'")
        super().__init__(initParamMap=initParamMap, **kwargs)
