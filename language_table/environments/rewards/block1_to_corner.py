"""Single-block debugging task: 'push <block> to the bottom left corner' (reference ``rewards/block1_to_corner.py``;
success within 0.08 m of a point 0.08 m inside the corner)."""
import enum

from language_table.environments import blocks as blocks_module
from language_table.environments.rewards import block2absolutelocation as b2a, task_info

BUFFER = 0.08
X_MIN, X_MAX, Y_MIN, Y_MAX = 0.15, 0.6, -0.3048, 0.3048
CENTER_X = (X_MAX - X_MIN) / 2.0 + X_MIN
CENTER_Y = (Y_MAX - Y_MIN) / 2.0 + Y_MIN
BLOCK2ABSOLUTELOCATION_TARGET_DISTANCE = 0.08


class Locations(enum.Enum):
    BOTTOM_LEFT = "bottom_left"


ABSOLUTE_LOCATIONS = {"bottom_left": [X_MAX - BUFFER, Y_MIN + BUFFER]}
LOCATION_SYNONYMS = {"bottom_left": ["bottom left of the board", "bottom left", "bottom left corner"]}
BLOCK2ABSOLUTELOCATION_VERBS = ["move the", "push the", "slide the"]


def generate_all_instructions(block_mode):
    return [f"{verb} {name} to the {syn}" for name in blocks_module.get_blocks_text_descriptions(block_mode)
            for loc in ABSOLUTE_LOCATIONS for syn in LOCATION_SYNONYMS[loc] for verb in BLOCK2ABSOLUTELOCATION_VERBS]


class Block1ToCornerLocationReward(b2a.BlockToAbsoluteLocationReward):
    _locations = ABSOLUTE_LOCATIONS
    _location_synonyms = LOCATION_SYNONYMS

    def _radius(self):
        return BLOCK2ABSOLUTELOCATION_TARGET_DISTANCE

    def reset(self, state, blocks_on_table):
        block = self._sample_object(blocks_on_table)
        location = self._rng.choice(sorted(self._locations.keys()))
        info = self.reset_to(state, block, location, blocks_on_table)
        return task_info.FAILURE if self._in_goal_region(state, block, self._target_translation) else info
