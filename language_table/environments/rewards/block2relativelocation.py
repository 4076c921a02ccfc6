"""'push <block> <direction> [slightly]' (reference ``rewards/block2relativelocation.py``): target = block position +
direction * {near 0.15 m, far 0.25 m}; success within 0.1 m."""
import numpy as np

from language_table.environments import blocks as blocks_module
from language_table.environments.rewards import reward as base_reward, synonyms, task_info

MAGNITUDES = {"near": 0.15, "far": 0.25}
UP, DOWN, LEFT, RIGHT = -1.0, 1.0, -1.0, 1.0
_R2 = float(np.sqrt(2.0))
DIRECTIONS = {
    "up": np.array([UP, 0.0]), "down": np.array([DOWN, 0.0]), "left": np.array([0.0, LEFT]), "right": np.array([0.0, RIGHT]),
    "diagonal_up_left": np.array([UP, LEFT]) / _R2, "diagonal_up_right": np.array([UP, RIGHT]) / _R2,
    "diagonal_down_left": np.array([DOWN, LEFT]) / _R2, "diagonal_down_right": np.array([DOWN, RIGHT]) / _R2,
}
BLOCK2RELATIVELOCATION_VERBS = ["move the", "push the", "slide the"]
SLIGHTLY_SYNONYMS = ["slightly", "a bit", "a little", "a little bit", "somewhat"]
DIRECTION_SYNONYMS = {"up": ["up", "upwards"], "down": ["down", "downwards"], "left": ["to the left", "left"],
                      "right": ["to the right", "right"]}
DIAGONAL_PREPOSITIONS = ["%s and %s", "%s and then %s", "diagonally %s and %s", "%s and %s diagonally"]
BLOCK2RELATIVELOCATION_TARGET_DISTANCE = 0.1


def enumerate_slightly_instruction(verb, block, direction):
    yield f"slightly {verb} {block} {direction}"
    for syn in SLIGHTLY_SYNONYMS:
        yield f"{verb} {block} {syn} {direction}"
        yield f"{verb} {block} {direction} {syn}"


def create_slightly_instruction(rng, verb, block, direction):
    mode = rng.choice(["slightly_first", "prefix", "suffix"])
    if mode == "slightly_first":
        return f"slightly {verb} {block} {direction}"
    syn = rng.choice(SLIGHTLY_SYNONYMS)
    return f"{verb} {block} {syn} {direction}" if mode == "prefix" else f"{verb} {block} {direction} {syn}"


def enumerate_diagonal_direction(direction):
    _, first, second = direction.split("_")
    for a in DIRECTION_SYNONYMS[first]:
        for b in DIRECTION_SYNONYMS[second]:
            for prep in DIAGONAL_PREPOSITIONS:
                yield prep % (a, b)


def get_diagonal_direction(rng, direction):
    _, first, second = direction.split("_")
    a, b = rng.choice(DIRECTION_SYNONYMS[first]), rng.choice(DIRECTION_SYNONYMS[second])
    return rng.choice(DIAGONAL_PREPOSITIONS) % (a, b)


def _direction_phrases(direction):
    return list(enumerate_diagonal_direction(direction)) if "diagonal" in direction else DIRECTION_SYNONYMS[direction]


def generate_all_instructions(block_mode):
    out = []
    for name in blocks_module.get_blocks_text_descriptions(block_mode):
        for verb in BLOCK2RELATIVELOCATION_VERBS:
            for direction in DIRECTIONS:
                for phrase in _direction_phrases(direction):
                    out.extend(enumerate_slightly_instruction(verb, name, phrase))
                    out.append(f"{verb} {name} {phrase}")
    return out


class BlockToRelativeLocationReward(base_reward.LanguageTableReward):
    def _sample_instruction(self, block, distance_mode, direction, blocks_on_table):
        verb = self._rng.choice(BLOCK2RELATIVELOCATION_VERBS)
        name = self._rng.choice(synonyms.get_block_synonyms(block, blocks_on_table))
        phrase = (get_diagonal_direction(self._rng, direction) if "diagonal" in direction
                  else self._rng.choice(DIRECTION_SYNONYMS[direction]))
        if distance_mode == "near":
            return create_slightly_instruction(self._rng, verb, name, phrase)
        return f"{verb} {name} {phrase}"

    def reset(self, state, blocks_on_table):
        for _ in range(101):
            self._block = self._sample_object(blocks_on_table)
            direction = self._rng.choice(sorted(DIRECTIONS.keys()))
            mode = self._rng.choice(sorted(MAGNITUDES.keys()))
            target = self._get_translation_for_block(self._block, state) + DIRECTIONS[direction] * MAGNITUDES[mode]
            if base_reward.target_inside_bounds(target):
                break
        else:
            return task_info.FAILURE
        self._instruction = self._sample_instruction(self._block, mode, direction, blocks_on_table)
        self._target_translation = np.copy(target)
        self._in_reward_zone_steps = 0
        return task_info.Block2RelativeLocationTaskInfo(instruction=self._instruction, block=self._block,
                                                        location=direction, target_translation=self._target_translation)

    def get_goal_region(self):
        return self._target_translation, BLOCK2RELATIVELOCATION_TARGET_DISTANCE

    def reward(self, state):
        d = np.linalg.norm(self._get_translation_for_block(self._block, state) - self._target_translation)
        return self._pay_if(d < BLOCK2RELATIVELOCATION_TARGET_DISTANCE)
