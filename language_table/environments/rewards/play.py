"""Open-ended 'play' instructions (reference ``rewards/play.py``): long-horizon arrangement goals with no programmatic
success test — the reward is always (0, False).  The instruction families (place-then-rest, k-then-rest, triangle,
orderings, colour groupings, lines, surround, sort) are generated combinatorially and sampled uniformly."""
import itertools

import numpy as np

from language_table.environments.rewards import reward as base_reward, task_info

BLOCKS4 = ["red moon", "blue cube", "green star", "yellow pentagon"]
BLOCKS8 = ["red moon", "red pentagon", "blue moon", "blue cube", "green cube", "green star", "yellow star", "yellow pentagon"]
LOCATIONS = ["top left corner", "top center", "top right corner", "center left", "center", "center right",
             "bottom left corner", "bottom center", "bottom right corner"]
ORDERINGS = list(itertools.permutations(BLOCKS4))
COLORS = ["red", "blue", "green", "yellow"]
SHAPES = ["moon", "cube", "star", "pentagon"]
NUMBERS = ["one", "two", "three", "four", "five", "six", "seven", "eight"]


def _pairs(locations):
    return [(a, b) for a in locations for b in locations if a != b]


def obj_in_place_then_remainder_in_other(blocks, locations):
    return ["put the %s in the %s, then put the rest of the blocks in the %s" % (b, l0, l1)
            for b in blocks for l0, l1 in _pairs(locations)]


def k_in_place_then_k_minus_1_in_other(blocks, locations):
    out = []
    for number in NUMBERS[:len(blocks)][:-1]:
        noun = "block" if number == "one" else "blocks"
        out += ["put %s %s in the %s, then put the rest in the %s" % (number, noun, l0, l1) for l0, l1 in _pairs(locations)]
    return out


def triangle_in_place_remainder_in_rest(locations):
    return ["make a triangle out of three blocks and put it in the %s of the board, then put the remainder in the %s of "
            "the board" % (l0, l1) for l0, l1 in _pairs(locations)]


def order_by(orderings):
    return ["order the blocks from %s: %s" % (o, ", ".join(seq)) for o in ("top to bottom", "left to right") for seq in orderings]


def colors_in_locations():
    return ["put the %s blocks in the %s" % (c, l) for c in COLORS for l in LOCATIONS]


def group_color_pairs():
    return ["group the blocks by color"] + ["put the %s blocks together and the %s blocks together" % (a, b)
                                            for a, b in itertools.combinations(COLORS, 2)]


def get_line_tasks():
    return ["make a %s line out of %s blocks" % (o, w) for o in ("vertical", "horizontal", "diagonal")
            for w in ["all the"] + COLORS + SHAPES]


def get_surround_tasks():
    return ["surround the %s with the other blocks" % b for b in BLOCKS8]


def get_shape_instructions():
    return ["make a %s out of all the blocks" % s for s in ("circle", "square", "triangle", "line", "smiley face", "letter L")]


def get_sort_tasks():
    return ["sort the blocks by color", "sort the blocks by shape", "put all the blocks in the center",
            "spread the blocks out evenly"]


def all_4block_instructions():
    return (obj_in_place_then_remainder_in_other(BLOCKS4, LOCATIONS) + k_in_place_then_k_minus_1_in_other(BLOCKS4, LOCATIONS) +
            triangle_in_place_remainder_in_rest(LOCATIONS) + order_by(ORDERINGS))


def all_8block_instructions():
    return (obj_in_place_then_remainder_in_other(BLOCKS8, LOCATIONS) + k_in_place_then_k_minus_1_in_other(BLOCKS8, LOCATIONS) +
            colors_in_locations() + group_color_pairs() + get_line_tasks() + get_surround_tasks() + get_shape_instructions() +
            get_sort_tasks())


def get_100_4block_instructions(num_train_per_family=20, num_test_per_family=5, return_train=True, seed=0):
    """A fixed, seeded sample of each 4-block family, split into train/test."""
    rng = np.random.RandomState(seed)
    train, test = [], []
    for family in (obj_in_place_then_remainder_in_other(BLOCKS4, LOCATIONS), k_in_place_then_k_minus_1_in_other(BLOCKS4, LOCATIONS),
                   triangle_in_place_remainder_in_rest(LOCATIONS), order_by(ORDERINGS)):
        family = list(family)
        rng.shuffle(family)
        train += family[:num_train_per_family]
        test += family[num_train_per_family:num_train_per_family + num_test_per_family]
    return train if return_train else test


def get_random_8block_instruction(rng):
    pool = all_8block_instructions()
    return pool[int(rng.randint(len(pool)))]


class PlayReward(base_reward.LanguageTableReward):
    def __init__(self, goal_reward, rng, delay_reward_steps, block_mode):
        super().__init__(goal_reward, rng, delay_reward_steps, block_mode)
        self._instruction = None

    def _sample_instruction(self, blocks_on_table):
        if len(blocks_on_table) <= 4:
            pool = get_100_4block_instructions()
            return pool[int(self._rng.randint(len(pool)))]
        return get_random_8block_instruction(self._rng)

    def reset(self, state, blocks_on_table):
        self._instruction = self._sample_instruction(blocks_on_table)
        self._in_reward_zone_steps = 0
        return task_info.PlayTaskInfo(instruction=self._instruction)

    def reward(self, state):
        return 0.0, False
