"""'push <block> to the <board location>' (reference ``rewards/block2absolutelocation.py``): nine named locations on
the board, success within 0.115 m (0.1 m for the centre)."""
import enum

import numpy as np

from language_table.environments import blocks as blocks_module
from language_table.environments.rewards import reward as base_reward, synonyms, task_info

X_BUFFER = 0.025
X_MIN_REAL, X_MAX_REAL = 0.15, 0.6                # the reachable board; the named locations sit X_BUFFER closer to the arm
Y_MIN_REAL, Y_MAX_REAL = -0.3048, 0.3048
X_MIN, X_MAX = X_MIN_REAL - X_BUFFER, X_MAX_REAL - X_BUFFER
Y_MIN, Y_MAX = Y_MIN_REAL, Y_MAX_REAL
CENTER_X = (X_MAX - X_MIN) / 2.0 + X_MIN
CENTER_Y = (Y_MAX - Y_MIN) / 2.0 + Y_MIN
BLOCK2ABSOLUTELOCATION_TARGET_DISTANCE = 0.115
BLOCK2ABSOLUTELOCATION_CENTER_TARGET_DISTANCE = 0.1


class Locations(enum.Enum):
    TOP = "top"
    TOP_LEFT = "top_left"
    TOP_RIGHT = "top_right"
    CENTER = "center"
    CENTER_LEFT = "center_left"
    CENTER_RIGHT = "center_right"
    BOTTOM = "bottom"
    BOTTOM_LEFT = "bottom_left"
    BOTTOM_RIGHT = "bottom_right"


_ROWS = {"top": X_MIN, "center": CENTER_X, "bottom": X_MAX}
_COLS = {"left": Y_MIN, "": CENTER_Y, "right": Y_MAX}
ABSOLUTE_LOCATIONS = {}
for _r, _x in _ROWS.items():
    for _c, _y in (("", CENTER_Y), ("left", Y_MIN), ("right", Y_MAX)):
        ABSOLUTE_LOCATIONS[_r if not _c else f"{_r}_{_c}"] = [_x, _y]
ABSOLUTE_LOCATIONS = {loc.value: ABSOLUTE_LOCATIONS[loc.value] for loc in Locations}

LOCATION_SYNONYMS = {
    "top": ["top side", "top", "towards your base"],
    "top_left": ["top left of the board", "top left", "upper left corner", "top left corner"],
    "top_right": ["top right of the board", "top right", "upper right corner", "top right corner"],
    "center": ["middle of the board", "center of the board", "center", "middle"],
    "center_left": ["left side of the board", "center left", "left side"],
    "center_right": ["right side of the board", "center right", "right side"],
    "bottom": ["bottom side", "bottom"],
    "bottom_left": ["bottom left of the board", "bottom left", "lower left corner", "bottom left corner"],
    "bottom_right": ["bottom right of the board", "bottom right", "lower right corner", "bottom right corner"],
}
BLOCK2ABSOLUTELOCATION_VERBS = ["move the", "push the", "slide the"]


def generate_all_instructions(block_mode):
    return [f"{verb} {name} to the {syn}" for name in blocks_module.get_blocks_text_descriptions(block_mode)
            for loc in ABSOLUTE_LOCATIONS for syn in LOCATION_SYNONYMS[loc] for verb in BLOCK2ABSOLUTELOCATION_VERBS]


class BlockToAbsoluteLocationReward(base_reward.LanguageTableReward):
    _locations = ABSOLUTE_LOCATIONS
    _location_synonyms = LOCATION_SYNONYMS

    def __init__(self, goal_reward, rng, delay_reward_steps, block_mode):
        super().__init__(goal_reward, rng, delay_reward_steps, block_mode)
        self._block = self._instruction = self._location = None

    def _radius(self):
        return (BLOCK2ABSOLUTELOCATION_CENTER_TARGET_DISTANCE if self._location == Locations.CENTER.value
                else BLOCK2ABSOLUTELOCATION_TARGET_DISTANCE)

    def _sample_instruction(self, block, blocks_on_table, location):
        verb = self._rng.choice(synonyms.PUSH_VERBS)
        name = self._rng.choice(synonyms.get_block_synonyms(block, blocks_on_table))
        return f"{verb} {name} to the {self._rng.choice(self._location_synonyms[location])}"

    def reset(self, state, blocks_on_table):
        block = self._sample_object(blocks_on_table)
        location = self._rng.choice(sorted(self._locations.keys()))
        info = self.reset_to(state, block, location, blocks_on_table)
        return task_info.FAILURE if self._in_goal_region(state, block, self._target_translation) else info

    def reset_to(self, state, block, location, blocks_on_table):
        self._block, self._location = block, location
        self._instruction = self._sample_instruction(block, blocks_on_table, location)
        self._target_translation = np.array(self._locations[location], dtype=np.float64)
        self._in_reward_zone_steps = 0
        return self.get_current_task_info(state)

    @property
    def target_translation(self):
        return self._target_translation

    def get_goal_region(self):
        return self._target_translation, self._radius()

    def _in_goal_region(self, state, block, target):
        return float(np.linalg.norm(self._get_translation_for_block(block, state) - np.array(target))) < self._radius()

    def reward_for(self, state, pushing_block, target_translation):
        return self._pay_if(self._in_goal_region(state, pushing_block, target_translation))

    def reward(self, state):
        return self.reward_for(state, self._block, self._target_translation)

    def reward_for_info(self, state, info):
        return self.reward_for(state, info.block, info.target_translation)

    def debug_info(self, state):
        return float(np.linalg.norm(self._get_translation_for_block(self._block, state) - self._target_translation))

    def get_current_task_info(self, state):
        return task_info.Block2LocationTaskInfo(instruction=self._instruction, block=self._block,
                                                location=self._location, target_translation=self._target_translation)
