"""Scripted pushing oracle (reference ``environments/oracles/oriented_push_oracle.py:44``): a small state machine —
move to a pre-push pose behind the block (opposite the goal), re-orient if the block drifted sideways, push towards
the goal — emitting clipped (dx, dy) effector actions."""
import numpy as np

from language_table.environments import constants


class OrientedPushOracle:
    def __init__(self, env, action_noise_std=0.0, max_step=0.03, rng=None):
        self._env = env
        self._noise = action_noise_std
        self._max_step = max_step
        self._rng = rng or np.random.RandomState(0)
        self.phase = "move_to_pre_block"

    def reset(self):
        self.phase = "move_to_pre_block"

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
