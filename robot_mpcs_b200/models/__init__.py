from robot_mpcs_b200.models.mpcBase import MpcBase, MpcConfiguration, RobotConfiguration
from robot_mpcs_b200.models.mpcModel import MpcModel
from robot_mpcs_b200.models.diff_drive_mpc_model import MpcDiffDriveModel
