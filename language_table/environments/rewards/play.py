"""Open-ended 'play' instructions (reference ``rewards/play.py``): long-horizon arrangement goals with no programmatic
success test — the reward is always ``(0, False)``.

Two vocabularies, as in the reference:

* ``BLOCK_4``: five families (lines / surround / corners / shapes; block-then-rest; k-then-rest; triangle-then-rest;
  orderings), each shuffled with ``random.seed(0)`` and cut to 20 training + 5 held-out instructions
  (``play.py:103-157``) — 100 training instructions in total.
* ``BLOCK_8``: eleven families (``play.py:172-316``); an instruction is drawn by first picking a family uniformly, then
  an instruction of it (``play.py:319-333``), so small families are not drowned by the 40 320 edge orderings.

``reset`` still samples a far-enough start / target pair and reports it in a ``Block2BlockTaskInfo`` (``play.py:359-391``):
the oracle and the dataset writers expect those two fields even though the instruction does not mention them.
"""
import itertools
import random

import numpy as np

from language_table.environments.rewards import constants, reward as base_reward, task_info

BLOCKS4 = ["red moon", "blue cube", "green star", "yellow pentagon"]
BLOCKS8 = ["red moon", "red pentagon", "blue moon", "blue cube", "green cube", "green star", "yellow star", "yellow pentagon"]
LOCATIONS = ["top left corner", "top center", "top right corner", "center left", "center", "center right",
             "bottom left corner", "bottom center", "bottom right corner"]
ORDERINGS = list(itertools.permutations(BLOCKS4))
COLORS = ["red", "blue", "green", "yellow"]
NUMBERS = ["one", "two", "three", "four", "five", "six", "seven", "eight"]
MAX_INSTRUCTION_LENGTH = 256


def _pairs(locations):
    return [(a, b) for a in locations for b in locations if a != b]


def _checked(inst):
    if len(inst) > MAX_INSTRUCTION_LENGTH:
        raise ValueError(f"Instruction greater than max length: {inst}")
    return inst


# ---------------------------------------------------------------------------------------------------- 4-block families
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


def expand_instructions(instructions):
    """``[(template, fillers or None)]`` → one string per filler (the bare template when there is none)."""
    out = []
    for template, fillers in instructions:
        out += [template] if fillers is None else [template % f for f in fillers]
    return out


_SIMPLE_4BLOCK = [
    ("put all the blocks in a line", None),
    ("put all the blocks in a %s line", ["horizontal", "vertical"]),
    ("put all the blocks in a vertical line on the %s side of the board", ["left", "center", "right"]),
    ("put all the blocks in a horizontal line on the %s side of the board", ["bottom", "center", "top"]),
    ("put the blocks in a diagonal line from the %s", ["top left to bottom right", "top right to bottom left"]),
    ("surround the %s with the other blocks", BLOCKS4),
    ("put all the blocks in the %s", LOCATIONS),
    ("put blocks in all four corners", None),
    ("make a %s shape out of the blocks", ["rectangle", "square", "diamond", "parallelogram"]),
]


def four_block_families():
    return [expand_instructions(_SIMPLE_4BLOCK), obj_in_place_then_remainder_in_other(BLOCKS4, LOCATIONS),
            k_in_place_then_k_minus_1_in_other(BLOCKS4, LOCATIONS), triangle_in_place_remainder_in_rest(LOCATIONS),
            order_by(ORDERINGS)]


def get_100_4block_instructions(num_train_per_family=20, num_test_per_family=5, return_train=True):
    """Per family: shuffle with Python's ``random`` seeded once with 0 (the reference's generator, so the split is the
    same list), keep ``num_train_per_family`` for training and the next ``num_test_per_family`` as held-out;
    ``num_train_per_family=0`` keeps everything for training."""
    gen = random.Random(0)
    train, test = [], []
    for family in four_block_families():
        gen.shuffle(family)
        if num_train_per_family:
            train += family[:num_train_per_family]
            test += family[num_train_per_family:num_train_per_family + num_test_per_family]
        else:
            train += family
    return train if return_train else test


# ---------------------------------------------------------------------------------------------------- 8-block families
def unique_color_combos():
    """Every colour pair (a, b) followed by the complementary pair (c, d): six tuples (each split appears twice, once per
    leading pair)."""
    pairs = list(itertools.combinations(COLORS, 2))
    out = []
    for a, b in pairs:
        c, d = next(p for p in pairs if a not in p and b not in p)
        out.append((a, b, c, d))
    return out


def colors_in_locations():
    return [_checked("put the %s blocks in the %s, the %s blocks in the %s, the %s blocks in the %s, and the %s blocks in the %s."
                     % (c[0], l[0], c[1], l[1], c[2], l[2], c[3], l[3]))
            for c in itertools.permutations(COLORS, 4) for l in itertools.permutations(LOCATIONS, 4)]


def group_color_pairs():
    return ["put the %s and %s blocks together in a group, then put the %s and %s blocks together in a group." % p
            for p in itertools.permutations(COLORS, 4)]


def group_color_pairs_in_locations():
    return ["put the %s and %s blocks together in the %s, then put the %s and %s blocks together in the %s."
            % (a, b, l0, c, d, l1) for a, b, c, d in unique_color_combos() for l0, l1 in itertools.permutations(LOCATIONS, 2)]


def get_colors_in_lines():
    modes = ("horizontal", "vertical")
    return ["make one %s line out of the %s and %s blocks, then make a %s line out of the %s and %s blocks"
            % (m0, a, b, m1, c, d) for m0 in modes for m1 in modes for a, b, c, d in unique_color_combos()]


def get_line_tasks():
    out = ["put the blocks in a line", "put all the blocks in a vertical line", "put all the blocks in a horizontal line"]
    out += ["put all the blocks in a vertical line on the %s of the board" % m for m in ("left", "center", "right")]
    out += ["put all the blocks in a horizontal line on the %s of the board" % m for m in ("bottom", "center", "top")]
    out += ["put the blocks in a diagonal line from the %s" % m for m in ("top left to bottom right", "top right to bottom left")]
    return out


def get_surround_tasks():
    return ["surround the %s with the others" % b for b in BLOCKS8]


_EDGE_LOCATIONS = ["top left", "top center", "top right", "center left", "center right", "bottom left", "bottom center",
                   "bottom right"]


def _edge_instruction(ordering):
    return _checked("put the " + "".join("%s to %s, " % bl for bl in zip(ordering, _EDGE_LOCATIONS)))


def blocks_in_order_outer_edge():
    """All 8! assignments of the blocks to the eight outer cells (40 320 strings; ``sample_outer_edge`` draws one
    without building the list)."""
    return [_edge_instruction(o) for o in itertools.permutations(BLOCKS8)]


def sample_outer_edge(rng):
    return _edge_instruction([BLOCKS8[i] for i in rng.permutation(len(BLOCKS8))])


def all_blocks_in_location():
    return ["put all the blocks in the %s" % l for l in LOCATIONS]


def k_blocks_in_location_i_rest_in_location_j():
    return ["put %d blocks in the %s, then the rest in the %s" % (k, l0, l1) for k in range(1, 8)
            for l0, l1 in itertools.permutations(LOCATIONS, 2)]


def get_shape_instructions():
    shapes = ["square", "triangle", "circle", "diamond", "parallelogram"] + list("GOLEATXVYUSCZNJ")
    # the doubled closing quote is in the reference's strings (``play.py:307``), hence in its datasets
    return (['make a "%s"" shape out of all the blocks' % s for s in shapes] +
            ["make a smiley face out of the blocks",
             "make a rainbow out of the blocks (red, yellow, green, blue in a semicircle)"])


def get_sort_tasks():
    return ["group the blocks by color"]


EIGHT_BLOCK_FAMILIES = [get_sort_tasks, colors_in_locations, group_color_pairs, get_colors_in_lines,
                        group_color_pairs_in_locations, get_line_tasks, get_surround_tasks, blocks_in_order_outer_edge,
                        all_blocks_in_location, k_blocks_in_location_i_rest_in_location_j, get_shape_instructions]
_FAMILY_CACHE = {}


def get_random_8block_instruction(rng):
    """Family first (uniform over the eleven), then an instruction of it.  The families are built once and cached; the
    two large ones (72 576 colour placements, 40 320 edge orderings) are sampled directly instead of enumerated."""
    fn = EIGHT_BLOCK_FAMILIES[int(rng.randint(len(EIGHT_BLOCK_FAMILIES)))]
    if fn is blocks_in_order_outer_edge:
        return sample_outer_edge(rng)
    if fn is colors_in_locations:
        c = [COLORS[i] for i in rng.permutation(4)]
        l = [LOCATIONS[i] for i in rng.permutation(len(LOCATIONS))[:4]]
        return ("put the %s blocks in the %s, the %s blocks in the %s, the %s blocks in the %s, and the %s blocks in the %s."
                % (c[0], l[0], c[1], l[1], c[2], l[2], c[3], l[3]))
    pool = _FAMILY_CACHE.get(fn)
    if pool is None:
        pool = _FAMILY_CACHE[fn] = fn()
    return pool[int(rng.randint(len(pool)))]


class PlayReward(base_reward.LanguageTableReward):
    """No success test: ``reward`` is always ``(0, False)``; episodes end on the step limit."""

    def __init__(self, goal_reward, rng, delay_reward_steps, block_mode):
        super().__init__(goal_reward, rng, delay_reward_steps, block_mode)
        self.block_mode = getattr(block_mode, "value", block_mode)
        self._all_instructions = get_100_4block_instructions(num_train_per_family=20) if self.block_mode == "BLOCK_4" else None
        self._instruction = self._start_block = self._target_block = None

    def _sample_instruction(self, start_block=None, target_block=None, blocks_on_table=None):
        if self.block_mode == "BLOCK_4":
            return self._all_instructions[int(self._rng.randint(len(self._all_instructions)))]
        if self.block_mode == "BLOCK_8":
            return get_random_8block_instruction(self._rng)
        raise ValueError("Unsupported block mode: %s" % self.block_mode)

    def reset(self, state, blocks_on_table):
        for _ in range(11):                                 # a start / target pair that is not already touching
            start, target = self._sample_objects(blocks_on_table)
            gap = np.linalg.norm(self._get_translation_for_block(start, state) - self._get_translation_for_block(target, state))
            if gap >= constants.TARGET_BLOCK_DISTANCE + 0.01:
                break
        else:
            return task_info.FAILURE
        self._start_block, self._target_block = start, target
        self._instruction = self._sample_instruction(start, target, blocks_on_table)
        self._in_reward_zone_steps = 0
        return task_info.Block2BlockTaskInfo(instruction=self._instruction, block1=start, block2=target)

    def reward(self, state):
        return 0.0, False
