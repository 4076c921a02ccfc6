"""The slice of the Bullet client interface the state helpers use, served by the kinematic board.

The reference keeps its scene inside PyBullet and reads / writes it through a client object (``ObjState.get_bullet_state``
/ ``set_bullet_state``, ``utils_pybullet.py:264-315``; ``LanguageTable.pybullet_client``, ``language_table.py:362``).
This environment has no physics engine: the scene is the tensors of :class:`rt1b200.envs.board.BatchedBoard` and the joint
vector of the kinematic xArm.  ``BoardClient`` answers the same calls from them, so code written against the reference's
save / restore API — or a user's own ``client.getBasePositionAndOrientation(block_id)`` — runs unchanged:

* body ids ``0 .. n_blocks-1`` are the blocks of ``blocks.ALL_BLOCKS`` order used by the environment, ``robot_id`` is the arm;
* a block that is not on the table reports the parking height ``z = -1`` (the reference parks unused blocks out of view);
* velocities are zero — the contact model is quasi-static — and ``resetBaseVelocity`` accepts and drops its arguments;
* joint info tuples carry index, name and type (revolute = 0) in Bullet's positions and placeholders elsewhere.
"""
from typing import List, Sequence, Tuple

import numpy as np
import torch
from scipy.spatial import transform

JOINT_REVOLUTE = 0
PARKING_Z = -1.0


class BoardClient:
    def __init__(self, env):
        self._env = env

    # ------------------------------------------------------------------------------------------------ bodies
    @property
    def robot_id(self) -> int:
        return len(self._env._all_blocks)

    def getNumBodies(self) -> int:
        return len(self._env._all_blocks) + (1 if self._env._robot is not None else 0)

    def _is_robot(self, body: int) -> bool:
        if body == self.robot_id and self._env._robot is not None:
            return True
        if not 0 <= body < len(self._env._all_blocks):
            raise ValueError(f"no body with id {body}")
        return False

    def getBodyInfo(self, body: int) -> Tuple[bytes, bytes]:
        name = "xarm6" if self._is_robot(body) else self._env._all_blocks[body]
        return b"base", name.encode()

    def getBasePositionAndOrientation(self, body: int):
        if self._is_robot(body):
            return (0.0, 0.0, 0.0), (0.0, 0.0, 0.0, 1.0)
        b = self._env._board
        x, y = (float(v) for v in b.block_xy[0, body])
        z = 0.0 if bool(b.block_mask[0, body]) else PARKING_Z
        quat = transform.Rotation.from_euler("z", float(b.block_yaw[0, body])).as_quat()
        return (x, y, z), tuple(float(v) for v in quat)

    def resetBasePositionAndOrientation(self, body: int, xyz: Sequence[float], quat: Sequence[float]) -> None:
        if self._is_robot(body):
            return                                              # the arm's base is bolted to the table
        b = self._env._board
        b.block_xy[0, body] = torch.tensor([float(xyz[0]), float(xyz[1])], dtype=b.block_xy.dtype)
        b.block_yaw[0, body] = float(transform.Rotation.from_quat(np.asarray(quat, float)).as_euler("xyz")[2])
        b.block_mask[0, body] = float(xyz[2]) > 0.5 * PARKING_Z

    def getBaseVelocity(self, body: int):
        self._is_robot(body)
        return (0.0, 0.0, 0.0), (0.0, 0.0, 0.0)

    def resetBaseVelocity(self, body: int, linearVelocity=None, angularVelocity=None) -> None:
        self._is_robot(body)

    # ------------------------------------------------------------------------------------------------ joints
    def getNumJoints(self, body: int) -> int:
        return len(self._env._robot.get_joint_positions()) if self._is_robot(body) else 0

    def _joint(self, body: int, joint: int) -> int:
        n = self.getNumJoints(body)
        if not 0 <= joint < n:
            raise ValueError(f"body {body} has {n} joints, asked for {joint}")
        return joint

    def getJointInfo(self, body: int, joint: int) -> tuple:
        j = self._joint(body, joint)
        from language_table.environments.utils import xarm_sim_robot as xr
        return (j, b"joint%d" % (j + 1), JOINT_REVOLUTE, 7 + j, 6 + j, 1, 0.0, 0.0, -xr.JOINT_LIMIT, xr.JOINT_LIMIT, 0.0,
                xr.MAX_JOINT_SPEED, b"link%d" % (j + 1), (0.0, 0.0, 1.0), (0.0, 0.0, 0.0), (0.0, 0.0, 0.0, 1.0), j - 1)

    def getJointState(self, body: int, joint: int) -> tuple:
        j = self._joint(body, joint)
        return float(self._env._robot.get_joint_positions()[j]), 0.0, (0.0,) * 6, 0.0

    def getJointStates(self, body: int, joints: Sequence[int]) -> List[tuple]:
        return [self.getJointState(body, j) for j in joints]

    def resetJointState(self, body: int, joint: int, targetValue: float, targetVelocity: float = 0.0) -> None:
        j = self._joint(body, joint)
        q = np.array(self._env._robot.get_joint_positions(), dtype=np.float64)
        q[j] = float(targetValue)
        self._env._robot.reset_joints(q)

    def stepSimulation(self) -> None:
        """One simulation sub-step of the arm servo (the board itself advances per control step in ``env.step``)."""
        if self._env._robot is not None:
            self._env._robot.step_simulation(1, dt=1.0 / self._env._step_frequency)
