"""xArm6 kinematics without a physics engine (reference ``environments/utils/xarm_sim_robot.py:40-220``).

The reference loads the xArm6 URDF into PyBullet and uses it for forward kinematics (``getLinkState``), inverse
kinematics (``calculateInverseKinematics``) and joint position control.  PyBullet is not available here, so this module
restates the arm as plain math:

* forward kinematics — the serial chain of the xArm6's six revolute joints (joint origins of the public xarm6 URDF:
  base→1 z 0.267; 1→2 roll -90 deg; 2→3 (0.0535, -0.2845, 0); 3→4 (0.0775, 0.3425, 0) roll -90 deg; 4→5 roll +90 deg;
  5→6 (0.076, 0.097, 0) roll -90 deg).  The reference's initial joint vector (``constants.INITIAL_JOINT_POSITIONS``)
  maps to translation (0.3006, -0.2003, 0.1452) with the tool pointing straight down, as ``constants.py`` states;
* ``forward_kinematics()`` reports the centre of mass of link 6, like ``getLinkState(...)[0]`` does (offset
  (0, 0.0006, -0.0095) m in the tool frame — the figure that reproduces the numbers of the reference's own
  ``xarm_sim_robot_test.py``: 0.714479 / -0.0006); ``forward_kinematics(link_frame=True)`` gives the URDF link frame;
* inverse kinematics — damped least squares on the 6-D pose error with a numerical Jacobian, warm-started from the
  current joints (the reference seeds Bullet's solver with the current joints as rest poses, ``:163-183``);
* position control — each simulation sub-step moves every joint a fraction of the way to its target, rate-limited, so the
  effector LAGS its target the way the motor-driven arm does between 10 Hz control steps (``language_table.py:599-646``).
"""
import math
from typing import Optional, Sequence

import numpy as np
from scipy.spatial import transform

from language_table.environments.utils.pose3d import Pose3d

# URDFs the reference loads into Bullet (``:24-36``); kept as names for callers, the kinematics below restate the xArm6 chain
XARM_URDF_PATH = "third_party/bullet/examples/pybullet/gym/pybullet_data/xarm/xarm6_robot.urdf"
XARM_WHITE_URDF_PATH = "third_party/bullet/examples/pybullet/gym/pybullet_data/xarm/xarm6_robot_white.urdf"
SUCTION_URDF_PATH = "third_party/py/language_table/environments/assets/suction/suction-head-long.urdf"
CYLINDER_URDF_PATH = "third_party/py/language_table/environments/assets/suction/cylinder.urdf"
CYLINDER_REAL_URDF_PATH = "third_party/py/language_table/environments/assets/suction/cylinder_real.urdf"
HOME_JOINT_POSITIONS = np.deg2rad([0, -20, -80, 0, 100, -30])

# (xyz, roll) of each joint's origin in its parent link; every joint turns about its own z axis
_CHAIN = (((0.0, 0.0, 0.267), 0.0), ((0.0, 0.0, 0.0), -math.pi / 2), ((0.0535, -0.2845, 0.0), 0.0),
          ((0.0775, 0.3425, 0.0), -math.pi / 2), ((0.0, 0.0, 0.0), math.pi / 2), ((0.076, 0.097, 0.0), -math.pi / 2))
LINK6_COM_OFFSET = np.array([0.0, 0.0006, -0.009521])
JOINT_LIMIT = 17.0                       # the reference passes +-17 rad to Bullet ("TODO: use real limits", :170-176)
MAX_JOINT_SPEED = 3.14                   # rad/s, the xArm6's rated joint speed
POSITION_GAIN = 0.2                      # fraction of the remaining joint error closed per simulation sub-step


def _rot_x(a: float) -> np.ndarray:
    c, s = math.cos(a), math.sin(a)
    return np.array([[1.0, 0.0, 0.0], [0.0, c, -s], [0.0, s, c]])


_FIXED = []
for _xyz, _roll in _CHAIN:
    _t = np.eye(4)
    _t[:3, :3] = _rot_x(_roll)
    _t[:3, 3] = _xyz
    _FIXED.append(_t)


def chain_forward(q: Sequence[float]) -> np.ndarray:
    """4x4 pose of link 6 in the base frame for joint angles ``q`` (radians)."""
    t = np.eye(4)
    for fixed, qi in zip(_FIXED, q):
        c, s = math.cos(qi), math.sin(qi)
        rz = np.array([[c, -s, 0.0, 0.0], [s, c, 0.0, 0.0], [0.0, 0.0, 1.0, 0.0], [0.0, 0.0, 0.0, 1.0]])
        t = t @ fixed @ rz
    return t


def _pose_error(t: np.ndarray, target_pos: np.ndarray, target_rot: np.ndarray) -> np.ndarray:
    """6-vector: position error and rotation-vector error (target * current^-1), both in the base frame."""
    dr = transform.Rotation.from_matrix(target_rot @ t[:3, :3].T).as_rotvec()
    return np.concatenate([target_pos - t[:3, 3], dr])


class XArmSimRobot:
    """Same public surface as the reference class, minus the Bullet client argument."""

    def __init__(self, pybullet_client=None, initial_joint_positions=HOME_JOINT_POSITIONS, end_effector: str = "none",
                 color: str = "default"):
        if end_effector not in ("none", "suction", "cylinder", "cylinder_real"):
            raise ValueError('end_effector "%s" is not supported.' % end_effector)
        if color not in ("default", "white"):
            raise ValueError("Unrecognized xarm color %s" % color)
        self.initial_joint_positions = np.asarray(initial_joint_positions, np.float64).copy()
        self.end_effector = end_effector if end_effector != "none" else None
        self.effector_link = 6
        self._q = self.initial_joint_positions.copy()
        self._q_target = self._q.copy()
        self._qd = np.zeros(6)

    # ------------------------------------------------------------------------------------------------ state
    def reset_joints(self, joint_values) -> None:
        """Teleport the joints (start of an episode); also becomes the control target."""
        self._q = np.asarray(joint_values, np.float64).copy()
        self._q_target = self._q.copy()
        self._qd = np.zeros(6)

    def get_joint_positions(self) -> np.ndarray:
        return self._q.copy()

    def get_joints_measured(self):
        return self._q.copy(), self._qd.copy(), np.zeros(6)

    # ------------------------------------------------------------------------------------------- kinematics
    def forward_kinematics(self, joint_positions: Optional[Sequence[float]] = None, link_frame: bool = False) -> Pose3d:
        t = chain_forward(self._q if joint_positions is None else joint_positions)
        pos = t[:3, 3] if link_frame else t[:3, 3] + t[:3, :3] @ LINK6_COM_OFFSET
        return Pose3d(translation=pos.copy(), rotation=transform.Rotation.from_matrix(t[:3, :3]))

    def inverse_kinematics(self, world_effector_pose: Pose3d, max_iterations: int = 100, residual_threshold: float = 1e-10,
                           damping: float = 1e-3) -> np.ndarray:
        """Joint angles that put link 6 (as reported by :meth:`forward_kinematics`) at the target pose."""
        target_rot = world_effector_pose.rotation.as_matrix()
        # the target refers to the reported point (link-6 centre of mass): move it back to the link frame
        target_pos = np.asarray(world_effector_pose.translation, np.float64) - target_rot @ LINK6_COM_OFFSET
        q = self._q.copy()
        eps = 1e-6
        for _ in range(max_iterations):
            t = chain_forward(q)
            err = _pose_error(t, target_pos, target_rot)
            if float(err @ err) < residual_threshold:
                break
            jac = np.empty((6, 6))
            for j in range(6):
                dq = q.copy()
                dq[j] += eps
                tj = chain_forward(dq)
                jac[:3, j] = (tj[:3, 3] - t[:3, 3]) / eps
                jac[3:, j] = transform.Rotation.from_matrix(tj[:3, :3] @ t[:3, :3].T).as_rotvec() / eps
            step = jac.T @ np.linalg.solve(jac @ jac.T + damping * np.eye(6), err)
            norm = float(np.linalg.norm(step))
            if norm > 1.0:                                  # trust region: at most one radian (in norm) per iteration
                step *= 1.0 / norm
            q = np.clip(q + step, -JOINT_LIMIT, JOINT_LIMIT)
        return q

    # ---------------------------------------------------------------------------------------------- control
    def set_target_effector_pose(self, world_effector_pose: Pose3d) -> None:
        self.set_target_joint_positions(self.inverse_kinematics(world_effector_pose))

    def set_target_joint_positions(self, target_joint_positions) -> None:
        self._q_target = np.asarray(target_joint_positions, np.float64).copy()

    def set_target_joint_velocities(self, target_joint_velocities) -> None:
        self._qd = np.clip(np.asarray(target_joint_velocities, np.float64), -MAX_JOINT_SPEED, MAX_JOINT_SPEED)
        self._q_target = None                                # velocity mode until a position target is set again

    def step_simulation(self, num_steps: int = 1, dt: float = 1.0 / 240.0) -> None:
        """Advance the joint servo by ``num_steps`` simulation sub-steps of ``dt`` seconds."""
        max_step = MAX_JOINT_SPEED * dt
        for _ in range(int(num_steps)):
            if self._q_target is None:
                self._q = np.clip(self._q + self._qd * dt, -JOINT_LIMIT, JOINT_LIMIT)
                continue
            delta = np.clip(POSITION_GAIN * (self._q_target - self._q), -max_step, max_step)
            self._qd = delta / dt
            self._q = self._q + delta

    def set_alpha_transparency(self, alpha: float) -> None:
        """Visual-only in the reference (GUI debugging); kept so callers need no special case."""
        self._alpha = float(alpha)
