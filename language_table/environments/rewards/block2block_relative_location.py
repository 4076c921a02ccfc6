"""'push <block> to the <side> of <block>' (reference ``rewards/block2block_relative_location.py``): the goal is an
invisible point 0.08 m (0.04 m per axis for diagonals) off the target block; success when the pushed block lies on the
segment from 50 % to 110 % of that offset (within 0.04 m) and the target block was not dragged more than 0.05 m."""
import enum
import itertools

import numpy as np

from language_table.environments import blocks as blocks_module
from language_table.environments.rewards import reward as base_reward, synonyms, task_info

MAGNITUDE_X = MAGNITUDE_Y = 0.08
MAGNITUDE_X_DIAG = MAGNITUDE_Y_DIAG = 0.04
BLOCK_DIAMETER = DELTA = 0.0075
DRAGGED_THRESHOLD = 0.05
UP, DOWN, LEFT, RIGHT = -1.0, 1.0, -1.0, 1.0
BLOCK2BLOCK_REL_LOCATION_TARGET_DISTANCE = 0.04


class Locations(enum.Enum):
    UP = "up"
    DOWN = "down"
    LEFT = "left"
    RIGHT = "right"
    DIAGONAL_UP_LEFT = "diagonal_up_left"
    DIAGONAL_UP_RIGHT = "diagonal_up_right"
    DIAGONAL_DOWN_LEFT = "diagonal_down_left"
    DIAGONAL_DOWN_RIGHT = "diagonal_down_right"


DIRECTIONS = {"up": [UP, 0.0], "down": [DOWN, 0.0], "left": [0.0, LEFT], "right": [0.0, RIGHT],
              "diagonal_up_left": [UP, LEFT], "diagonal_up_right": [UP, RIGHT],
              "diagonal_down_left": [DOWN, LEFT], "diagonal_down_right": [DOWN, RIGHT]}
DIRECTION_IDS = list(DIRECTIONS.keys())


def get_unique_task_strings():
    """``"<block>-<target block>-<direction>"`` → task id, ids in sorted-string order (reference ``:84-98``)."""
    return {s: i for i, s in enumerate(sorted(
        f"{a}-{b}-{d}" for a in blocks_module.ALL_BLOCKS for b in blocks_module.ALL_BLOCKS for d in DIRECTIONS))}


UNIQUE_TASK_STRINGS = get_unique_task_strings()
NUM_UNIQUE_TASKS = len(UNIQUE_TASK_STRINGS)

BLOCK2BLOCKRELATIVELOCATION_VERBS = ["move the", "push the", "put the", "bring the", "slide the"]
DIRECTION_SYNONYMS = {
    "up": ["above the", "to the top side of the", "to the top of the"],
    "down": ["below the", "to the bottom side of the", "to the bottom of the"],
    "left": ["just left of the", "to the left of the", "left of the", "to the left side of the"],
    "right": ["just right of the", "to the right of the", "right of the", "to the right side of the"],
    "diagonal_up_left": ["to the top left side of the", "to the top left of the", "diagonally up and to the left of the"],
    "diagonal_up_right": ["to the top right side of the", "to the top right of the", "diagonally up and to the right of the"],
    "diagonal_down_left": ["to the bottom left side of the", "to the bottom left of the", "diagonally down and to the left of the"],
    "diagonal_down_right": ["to the bottom right side of the", "to the bottom right of the", "diagonally down and to the right of the"],
}


def _offset(direction, scale=1.0):
    diag = "diagonal" in direction
    mag = np.array([MAGNITUDE_X_DIAG if diag else MAGNITUDE_X, MAGNITUDE_Y_DIAG if diag else MAGNITUDE_Y]) * scale
    return np.array(DIRECTIONS[direction]) * mag


def is_block2block_relative_pair(xy_block, xy_target):
    return any(np.linalg.norm(np.array(xy_block) + _offset(d) - xy_target) < 1e-6 for d in DIRECTIONS)


def generate_all_instructions(block_mode):
    names = blocks_module.get_blocks_text_descriptions(block_mode)
    return [f"{verb} {a} {syn} {b}" for a, b in itertools.permutations(names, 2)
            for verb in BLOCK2BLOCKRELATIVELOCATION_VERBS for d in DIRECTIONS for syn in DIRECTION_SYNONYMS[d]]


class BlockToBlockRelativeLocationReward(base_reward.LanguageTableReward):
    def __init__(self, goal_reward, rng, delay_reward_steps, block_mode):
        super().__init__(goal_reward, rng, delay_reward_steps, block_mode)
        self._target_block = self._block = self._direction = self._instruction = None

    def target_translation_for(self, state, target_block, direction, scale=1.0):
        return self._get_translation_for_block(target_block, state) + _offset(direction, scale)

    def get_current_task_info(self, state):
        if self._target_block is None:
            raise ValueError("must call .reset first")
        self._target_translation = self.target_translation_for(state, self._target_block, self._direction)
        return task_info.Block2BlockRelativeLocationTaskInfo(
            instruction=self._instruction, block=self._block, target_translation=self._target_translation,
            target_block=self._target_block, direction=self._direction)

    def _sample_instruction(self, block, target_block, direction, blocks_on_table):
        verb = self._rng.choice(synonyms.PUSH_VERBS)
        a = self._rng.choice(synonyms.get_block_synonyms(block, blocks_on_table))
        b = self._rng.choice(synonyms.get_block_synonyms(target_block, blocks_on_table))
        return f"{verb} {a} {self._rng.choice(DIRECTION_SYNONYMS[direction])} {b}"

    instruction_for = _sample_instruction

    def instruction_for(self, pushing_block, target_block, direction, blocks_on_table):
        """A freshly sampled phrasing of a given (block, target block, direction) task (reference ``:274-277``)."""
        return self._sample_instruction(pushing_block, target_block, direction, blocks_on_table)

    def reset(self, state, blocks_on_table):
        for _ in range(101):
            block, target_block = self._sample_objects(blocks_on_table)
            direction = self._rng.choice(list(DIRECTIONS.keys()))
            if base_reward.target_inside_bounds(self.target_translation_for(state, target_block, direction)):
                break
        else:
            return task_info.FAILURE
        info = self.reset_to(state, block, target_block, direction, blocks_on_table)
        self._in_reward_zone_steps = 0
        if self.reward_for(state, block, target_block, direction, delay_reward_steps=0)[1]:
            return task_info.FAILURE
        return info

    @property
    def target_translation(self):
        return self._target_translation

    def reset_to(self, state, block, target_block, direction, blocks_on_table):
        self._block, self._target_block, self._direction = block, target_block, direction
        self._target_block_reset_translation = np.copy(self._get_translation_for_block(target_block, state))
        self._target_translation = self.target_translation_for(state, target_block, direction)
        self._instruction = self._sample_instruction(block, target_block, direction, blocks_on_table)
        return self.get_current_task_info(state)

    def get_distance_for(self, state, pushing_block, target_block, direction):
        return float(np.linalg.norm(self._get_translation_for_block(pushing_block, state) -
                                    self.target_translation_for(state, target_block, direction)))

    def debug_info(self, state):
        return self.get_distance_for(state, self._block, self._target_block, self._direction)

    def reward_for(self, state, pushing_block, target_block, direction, delay_reward_steps):
        pusher = self._get_translation_for_block(pushing_block, state)
        anchor = self._get_translation_for_block(target_block, state)
        diff = self.target_translation_for(state, target_block, direction) - anchor
        pts = anchor + np.linspace(diff * 0.5, diff * 1.1, 10)
        on_line = bool((np.linalg.norm(pts - pusher, axis=1) < BLOCK2BLOCK_REL_LOCATION_TARGET_DISTANCE).any())
        dragged = np.linalg.norm(self._target_block_reset_translation - anchor) > DRAGGED_THRESHOLD
        return self._pay_if(on_line and not dragged, delay_reward_steps)

    def reward(self, state):
        return self.reward_for(state, self._block, self._target_block, self._direction, self._delay_reward_steps)

    def get_goal_region(self):
        return self._target_translation, BLOCK2BLOCK_REL_LOCATION_TARGET_DISTANCE

    def reward_for_info(self, state, info):
        return self.reward_for(state, info.block, info.target_block, info.direction, self._delay_reward_steps)

    def get_current_task_id(self):
        return UNIQUE_TASK_STRINGS[f"{self._block}-{self._target_block}-{self._direction}"]
