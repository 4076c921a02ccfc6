"""'point at <block>' (reference ``rewards/point2block.py:37-98``; success: effector target within 0.05 m of the block)."""
import numpy as np

from language_table.environments import blocks as blocks_module
from language_table.environments.rewards import constants, reward as base_reward, synonyms, task_info


def generate_all_instructions(block_mode):
    return [f"{prep} {name}" for name in blocks_module.get_blocks_text_descriptions(block_mode)
            for prep in synonyms.POINT_PREPOSITIONS]


class PointToBlockReward(base_reward.LanguageTableReward):
    def _dist(self, block, state):
        return float(np.linalg.norm(self._get_translation_for_block(block, state) -
                                    np.array(state["effector_target_translation"])))

    def reset(self, state, blocks_on_table):
        for _ in range(11):
            block = self._sample_object(blocks_on_table)
            if self._dist(block, state) >= constants.TARGET_BLOCK_DISTANCE + 0.01:
                break
        else:
            return task_info.FAILURE
        self._block = block
        name = self._rng.choice(synonyms.get_block_synonyms(block, blocks_on_table))
        self._instruction = f"{self._rng.choice(synonyms.POINT_PREPOSITIONS)} {name}"
        self._in_reward_zone_steps = 0
        return task_info.Point2BlockTaskInfo(instruction=self._instruction, block_target=block)

    def reward(self, state):
        return self._pay_if(self._dist(self._block, state) < constants.TARGET_BLOCK_DISTANCE)
