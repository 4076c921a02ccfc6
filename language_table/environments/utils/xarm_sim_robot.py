"""Kinematic stand-in for ``XArmSimRobot`` (reference ``environments/utils/xarm_sim_robot.py:40-220``).

The reference drives a 6-DoF xArm URDF through PyBullet IK + position control just to move a cylinder in the plane.
Without a physics engine the arm itself is not simulated: this class keeps the effector pose and reproduces the
interface the environment and oracles use (forward kinematics = stored pose, inverse kinematics = identity on the
planar target, position control = move to target)."""
import numpy as np
from scipy.spatial import transform

from language_table.environments import constants
from language_table.environments.utils.pose3d import Pose3d

HOME_JOINT_POSITIONS = np.deg2rad([0, -20, -80, 0, 100, -30])


class XArmSimRobot:
    def __init__(self, initial_joint_positions=HOME_JOINT_POSITIONS, end_effector="none", color="default"):
        self.initial_joint_positions = np.asarray(initial_joint_positions, np.float64)
        self._pose = Pose3d(rotation=transform.Rotation.from_rotvec([0, np.pi, 0]),
                            translation=np.array([0.3, -0.2, constants.EFFECTOR_HEIGHT]))
        self._target = self._pose

    def reset_joints(self, joint_values=None):
        self._target = self._pose

    def forward_kinematics(self) -> Pose3d:
        return self._pose

    def inverse_kinematics(self, world_effector_pose: Pose3d, max_iterations=100, residual_threshold=1e-10):
        t = np.asarray(world_effector_pose.translation, np.float64)
        return np.array([t[0], t[1], t[2], 0.0, 0.0, 0.0])       # "joint values" of the planar stand-in

    def set_target_effector_pose(self, world_effector_pose: Pose3d):
        self._target = world_effector_pose
        self.set_target_joint_positions(self.inverse_kinematics(world_effector_pose))

    def set_target_joint_positions(self, target_joint_positions):
        t = np.asarray(target_joint_positions, np.float64)
        self._pose = Pose3d(rotation=self._target.rotation, translation=np.array([t[0], t[1], t[2]]))

    def get_joint_positions(self):
        t = self._pose.translation
        return np.array([t[0], t[1], t[2], 0.0, 0.0, 0.0])
