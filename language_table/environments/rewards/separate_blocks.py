"""'separate <block> from <blocks>' (reference ``rewards/separate_blocks.py``): pick the block whose neighbours (< 0.08 m)
are on average closest, the goal is 0.1 m away from the neighbours' centroid on the far side; success within 0.025 m."""
import numpy as np

from language_table.environments import blocks as blocks_module
from language_table.environments.rewards import reward as base_reward, task_info

CONSIDERED_JOINED_THRESHOLD = 0.08
MAGNITUDE = 0.1
DISTANCE_TO_TARGET_THRESHOLD = 0.025
SEPARATE_TEMPLATES = ["pull the %s apart from the %s", "move the %s away from the %s", "separate the %s from the %s"]
GROUP_SYNONYMS = ["group", "clump", "group of blocks"]
REST = "rest of the blocks"


def _avoid_phrase(avoid, n_on_table, group_syn, rng=None):
    phrase = None
    if len(avoid) == n_on_table - 1:
        phrase = REST
    if len(avoid) == 1:
        phrase = avoid[0]
    if len(avoid) == 2:
        phrase = "%s and %s" % tuple(avoid)
    if len(avoid) == 3:
        specific = "%s, %s, and %s" % tuple(avoid)
        phrase = specific if rng is None else rng.choice([specific, group_syn])
    if len(avoid) >= 4:
        phrase = group_syn
    return phrase


def generate_all_instructions(block_mode):
    names = blocks_module.get_blocks_text_descriptions(block_mode)
    out = []
    for name in names:
        for idx in range(1, len(names)):
            for group_syn in GROUP_SYNONYMS:
                phrase = _avoid_phrase(names[:idx], len(names), group_syn)
                out.extend(t % (name, phrase) for t in SEPARATE_TEMPLATES)
    return out


class SeparateBlocksReward(base_reward.LanguageTableReward):
    def __init__(self, goal_reward, rng, delay_reward_steps, block_mode):
        super().__init__(goal_reward, rng, delay_reward_steps, block_mode)
        self._instruction = self._block = self._avoid_blocks = self._avoid_centroid_xy = None

    def get_current_task_info(self, state):
        if self._block is None:
            raise ValueError("must call .reset first")
        self._target_translation = self.target_translation_for(state, self._block, self._avoid_blocks)
        return task_info.SeparateBlocksTaskInfo(instruction=self._instruction, block=self._block,
                                                avoid_blocks=self._avoid_blocks, target_translation=self._target_translation)

    def _sample_instruction(self, block, avoid_blocks, blocks_on_table):
        name = self._get_block_synonym(block, blocks_on_table)
        avoid = [self._get_block_synonym(b, blocks_on_table) for b in avoid_blocks]
        phrase = _avoid_phrase(avoid, len(blocks_on_table), self._rng.choice(GROUP_SYNONYMS), self._rng)
        return self._rng.choice(SEPARATE_TEMPLATES) % (name, phrase)

    def _get_blocks_to_separate(self, state, blocks_on_table):
        xy = {b: self._get_translation_for_block(b, state) for b in blocks_on_table}
        best = None
        for src in blocks_on_table:
            near = sorted(((b, float(np.linalg.norm(xy[src] - xy[b]))) for b in blocks_on_table if b != src), key=lambda t: t[1])
            near = [t for t in near if t[1] < CONSIDERED_JOINED_THRESHOLD]
            avg = float(np.mean([t[1] for t in near])) if near else np.inf
            if best is None or avg < best[2]:
                best = (src, [t[0] for t in near], avg)
        return best

    def _get_avoid_direction(self, state, push_block, avoid_blocks):
        centroid = np.mean([self._get_translation_for_block(b, state) for b in avoid_blocks], axis=0)
        self._avoid_centroid_xy = centroid
        to_centroid = centroid - self._get_translation_for_block(push_block, state)
        return -to_centroid / (np.linalg.norm(to_centroid) + np.finfo(np.float32).eps)

    def target_translation_for(self, state, block, avoid_blocks):
        direction = self._get_avoid_direction(state, block, avoid_blocks)
        return self._avoid_centroid_xy + direction * MAGNITUDE

    def reset(self, state, blocks_on_table):
        push_block, avoid_blocks, _ = self._get_blocks_to_separate(state, blocks_on_table)
        if not avoid_blocks:
            return task_info.FAILURE
        if not base_reward.target_inside_bounds(self.target_translation_for(state, push_block, avoid_blocks)):
            return task_info.FAILURE        # deterministic given the state: retrying (as upstream does 100x) cannot help
        return self.reset_to(state, push_block, avoid_blocks, blocks_on_table)

    @property
    def target_translation(self):
        return self._target_translation

    def reset_to(self, state, block, avoid_blocks, blocks_on_table):
        self._block, self._avoid_blocks = block, avoid_blocks
        self._target_translation = self.target_translation_for(state, block, avoid_blocks)
        self._instruction = self._sample_instruction(block, avoid_blocks, blocks_on_table)
        self._in_reward_zone_steps = 0
        return self.get_current_task_info(state)

    def _dist_for(self, state, push_block, target_translation):
        return float(np.linalg.norm(self._get_translation_for_block(push_block, state) - target_translation))

    def debug_info(self, state):
        return self._dist_for(state, self._block, self._target_translation)

    def reward_for(self, state, push_block, target_translation):
        return self._pay_if(self._dist_for(state, push_block, target_translation) < DISTANCE_TO_TARGET_THRESHOLD)

    def reward(self, state):
        return self.reward_for(state, self._block, self._target_translation)

    def reward_for_info(self, state, info):
        return self.reward_for(state, info.block, info.target_translation)

    def get_goal_region(self):
        return self._target_translation, DISTANCE_TO_TARGET_THRESHOLD
