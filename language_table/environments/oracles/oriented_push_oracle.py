"""Scripted pushing oracle (reference ``environments/oracles/oriented_push_oracle.py:44``): a small state machine —
move to a pre-push pose behind the block (opposite the goal), re-orient if the block drifted sideways, push towards
the goal — emitting clipped (dx, dy) effector actions."""
import dataclasses
from typing import Any

import numpy as np

from language_table.environments import constants


@dataclasses.dataclass
class PushingInfo:
    """The per-step quantities of the push state machine (same fields as the reference's record,
    ``oriented_push_oracle.py:27-41``); ``OrientedPushOracle.get_pushing_info`` fills it."""
    xy_block: Any = None
    xy_ee: Any = None
    xy_pre_block: Any = None
    xy_dir_block_to_target: Any = None
    xy_delta_to_nexttoblock: Any = None
    xy_delta_to_touchingblock: Any = None
    xy_dir_block_to_ee: Any = None
    theta_threshold_to_orient: Any = None
    theta_threshold_flat_enough: Any = None
    theta_error: Any = None
    obstacle_poses: Any = None
    distance_to_target: Any = None


class OrientedPushOracle:
    def __init__(self, env, action_noise_std=0.0, max_step=0.03, rng=None):
        self._env = env
        self._noise = action_noise_std
        self._max_step = max_step
        self._rng = rng or np.random.RandomState(0)
        self.phase = "move_to_pre_block"

    def reset(self):
        self.phase = "move_to_pre_block"

    # planar helpers of the reference's oracle (``oriented_push_oracle.py:58-68``)
    def get_theta_from_vector(self, vector):
        return np.arctan2(vector[1], vector[0])

    def theta_to_rotation2d(self, theta):
        c, s = np.cos(theta), np.sin(theta)
        return np.array([[c, -s], [s, c]])

    def rotate(self, theta, xy_dir_block_to_ee):
        return self.theta_to_rotation2d(theta) @ np.asarray(xy_dir_block_to_ee)

    def get_pushing_info(self, state=None) -> PushingInfo:
        """Geometry of the current step: block / effector / pre-push positions, the push direction, the signed angle
        between "block → effector" and "target → block" (zero when the effector sits exactly behind the block), the
        other blocks as obstacles and the remaining distance."""
        state = state or self._env.compute_state()
        eff = np.asarray(state["effector_target_translation"], float)
        block, goal = self._goal_and_block(state)
        to_goal = goal - block
        dist = float(np.linalg.norm(to_goal))
        direction = to_goal / max(dist, 1e-6)
        rel = eff - block
        dir_ee = rel / max(float(np.linalg.norm(rel)), 1e-6)
        near = constants.BLOCK_RADIUS + constants.EFFECTOR_RADIUS
        theta_error = self.get_theta_from_vector(dir_ee) - self.get_theta_from_vector(-direction)
        theta_error = (theta_error + np.pi) % (2 * np.pi) - np.pi
        name = self._block_name()
        obstacles = [np.asarray(v, float) for k, v in state.items()
                     if k.startswith("block_") and k.endswith("_translation") and k != "block_%s_translation" % name]
        return PushingInfo(xy_block=block, xy_ee=eff, xy_pre_block=block - direction * (near + 0.015),
                           xy_dir_block_to_target=direction, xy_delta_to_nexttoblock=block + dir_ee * (near + 0.015) - eff,
                           xy_delta_to_touchingblock=block + dir_ee * near - eff, xy_dir_block_to_ee=dir_ee,
                           theta_threshold_to_orient=0.2, theta_threshold_flat_enough=0.03, theta_error=float(theta_error),
                           obstacle_poses=obstacles, distance_to_target=dist)

    def _block_name(self):
        info = getattr(self._env, "_task_info", None)
        return getattr(info, "block1", None) or getattr(info, "block", None) or getattr(info, "block_target", None)

    def _goal_and_block(self, state):
        info = getattr(self._env, "_task_info", None)
        block = getattr(info, "block1", None) or getattr(info, "block", None) or getattr(info, "block_target", None)
        if hasattr(info, "block2"):
            goal = np.asarray(state["block_%s_translation" % info.block2], float)
        elif hasattr(info, "target_translation") and info.target_translation is not None:
            goal = np.asarray(info.target_translation, float)
        else:
            goal = np.asarray(state["block_%s_translation" % block], float)
        return np.asarray(state["block_%s_translation" % block], float), goal

    def _clip(self, delta):
        n = np.linalg.norm(delta)
        if n > self._max_step:
            delta = delta / n * self._max_step
        if self._noise > 0:
            delta = delta + self._rng.normal(0, self._noise, 2)
        return delta.astype(np.float32)

    def action(self, state=None):
        state = state or self._env.compute_state()
        eff = np.asarray(state["effector_target_translation"], float)
        block, goal = self._goal_and_block(state)
        to_goal = goal - block
        dist = np.linalg.norm(to_goal)
        direction = to_goal / max(dist, 1e-6)
        standoff = constants.BLOCK_RADIUS + constants.EFFECTOR_RADIUS + 0.015
        pre = block - direction * standoff
        rel = eff - block
        along, lateral = float(rel @ direction), rel - (rel @ direction) * direction
        if self.phase == "move_to_pre_block":
            if along > -0.01 and np.linalg.norm(lateral) < 0.06:          # in front of the block: swing around it
                side = np.array([-direction[1], direction[0]]) * (1.0 if lateral @ np.array([-direction[1], direction[0]]) >= 0 else -1.0)
                return self._clip(block + side * 0.07 - direction * 0.02 - eff)
            if np.linalg.norm(pre - eff) < 0.01:
                self.phase = "push_block"
            return self._clip(pre - eff)
        if np.linalg.norm(lateral) > 0.015 or along > 0:                     # lost the block: line up again
            self.phase = "move_to_pre_block"
            return self._clip(pre - eff)
        return self._clip(direction * min(self._max_step, dist + 0.02))
