"""Base class of the sparse board rewards (reference ``rewards/reward.py:24-74``).

Every task family follows one protocol: ``reset(state, blocks_on_table)`` samples a task and returns its task-info
(or ``task_info.FAILURE``), ``reward(state)`` returns ``(reward, done)`` where the goal reward is paid once the
success predicate has held for ``delay_reward_steps`` consecutive calls.
"""
import numpy as np

from language_table.environments import constants


class LanguageTableReward:
    def __init__(self, goal_reward, rng, delay_reward_steps, block_mode):
        self._block_mode = block_mode
        self._goal_reward = goal_reward
        self._rng = rng
        self._delay_reward_steps = delay_reward_steps
        self._in_reward_zone_steps = None
        self._target_translation = None

    def seed(self, rng):
        self._rng = rng

    def get_goal_region(self):
        return None, None

    # -- helpers shared by the families --------------------------------------------------------------
    def _pay_if(self, in_zone: bool, delay=None):
        """Sparse reward with the 'stay in the zone for N steps' rule."""
        delay = self._delay_reward_steps if delay is None else delay
        if not in_zone:
            return 0.0, False
        if self._in_reward_zone_steps >= delay:
            return self._goal_reward, True
        self._in_reward_zone_steps += 1
        return 0.0, False

    def _get_block_synonym(self, block, blocks_on_table):
        from language_table.environments.rewards import synonyms
        return self._rng.choice(synonyms.get_block_synonyms(block, blocks_on_table))

    def _get_pose_for_block(self, block, state):
        return state["block_%s_translation" % block], state["block_%s_orientation" % block]

    def _get_translation_for_block(self, block, state):
        return np.array(self._get_pose_for_block(block, state)[0])

    def _sample_object(self, blocks_on_table):
        return self._rng.choice(blocks_on_table)

    def _sample_objects(self, blocks_on_table):
        a, b = self._rng.choice(blocks_on_table, 2, replace=False)
        return a, b


def target_inside_bounds(target, buffer=constants.WORKSPACE_BOUNDS_BUFFER):
    x, y = target
    return (constants.X_MIN + buffer < x < constants.X_MAX - buffer and
            constants.Y_MIN + buffer < y < constants.Y_MAX - buffer)
