"""'push <block> to <block>' (reference ``rewards/block2block.py:43-118``; success: centres closer than 0.05 m)."""
import itertools

import numpy as np

from language_table.environments import blocks as blocks_module
from language_table.environments.rewards import constants, reward as base_reward, synonyms, task_info


def generate_all_instructions(block_mode):
    names = blocks_module.get_blocks_text_descriptions(block_mode)
    return [f"{verb} {a} {prep} {b}" for a, b in itertools.permutations(names, 2)
            for verb in synonyms.PUSH_VERBS for prep in synonyms.PREPOSITIONS]


class BlockToBlockReward(base_reward.LanguageTableReward):
    def _sample_instruction(self, start_block, target_block, blocks_on_table):
        verb = self._rng.choice(synonyms.PUSH_VERBS)
        a = self._rng.choice(synonyms.get_block_synonyms(start_block, blocks_on_table))
        b = self._rng.choice(synonyms.get_block_synonyms(target_block, blocks_on_table))
        prep = self._rng.choice(synonyms.PREPOSITIONS)
        return f"{verb} {a} {prep} {b}"

    def _distance(self, state):
        a = self._get_translation_for_block(self._start_block, state)
        b = self._get_translation_for_block(self._target_block, state)
        return float(np.linalg.norm(a - b)), b

    def reset(self, state, blocks_on_table):
        for _ in range(11):
            start, target = self._sample_objects(blocks_on_table)
            d = np.linalg.norm(self._get_translation_for_block(start, state) - self._get_translation_for_block(target, state))
            if d >= constants.TARGET_BLOCK_DISTANCE + 0.01:
                break
        else:
            return task_info.FAILURE
        self._start_block, self._target_block = start, target
        self._instruction = self._sample_instruction(start, target, blocks_on_table)
        self._in_reward_zone_steps = 0
        return task_info.Block2BlockTaskInfo(instruction=self._instruction, block1=start, block2=target)

    def get_goal_region(self):
        return self._target_translation, constants.TARGET_BLOCK_DISTANCE

    def reward(self, state):
        dist, self._target_translation = self._distance(state)
        return self._pay_if(dist < constants.TARGET_BLOCK_DISTANCE)
