"""Block sets of the Language-Table board (reference ``language_table/environments/blocks.py:24-160``): four colours x
four shapes, the fixed 1/4/8-block boards, the pole variants and the seeded train/test split of all 4..10-block
combinations."""
import enum
import collections
import itertools

import numpy as np


class LanguageTableBlockVariants(enum.Enum):
    BLOCK_1 = "BLOCK_1"
    BLOCK_4 = "BLOCK_4"
    BLOCK_8 = "BLOCK_8"
    BLOCK_4_WPOLE = "BLOCK_4_WPOLE"
    BLOCK_8_WPOLE = "BLOCK_8_WPOLE"
    N_CHOOSE_K = "N_CHOOSE_K"


BLOCK_VARIANTS = [v.value for v in LanguageTableBlockVariants]
COLORS = ["red", "blue", "green", "yellow"]
SHAPES = ["moon", "cube", "star", "pentagon"]
ALL_BLOCKS = ["_".join(cs) for cs in itertools.product(COLORS, SHAPES)]
POLE = "purple_pole"
DUMMY_START_BLOCK = "red_moon"
# URDF of every block in the reference's path convention (its ``blocks.py:85-112``; resolved against the local asset directory
# by ``utils_pybullet.resolve_asset_path`` / ``rt1b200.envs.assets``), red / blue / yellow / green x moon / cube / star / pentagon
_ASSET_PREFIX = "third_party/py/language_table/environments/assets/"
BLOCK_URDF_PATHS = collections.OrderedDict(
    (f"{c}_{sh}", f"{_ASSET_PREFIX}blocks/{c}_{sh}.urdf") for c in ("red", "blue", "yellow", "green")
    for sh in ("moon", "cube", "star", "pentagon"))
POLE_URDF_PATHS = collections.OrderedDict(purple_pole=f"{_ASSET_PREFIX}blocks/purple_pole.urdf")

FIXED_1_COMBINATION = ["green_star"]
FIXED_4_COMBINATION = ("red_moon", "blue_cube", "green_star", "yellow_pentagon")
FIXED_8_COMBINATION = ("red_moon", "red_pentagon", "blue_moon", "blue_cube", "green_cube", "green_star", "yellow_star",
                       "yellow_pentagon")
FIXED_4_COMBINATION_WPOLE = FIXED_4_COMBINATION + (POLE,)
FIXED_8_COMBINATION_WPOLE = FIXED_8_COMBINATION + (POLE,)

MIN_K, MAX_K = 4, 10
_all = [c for k in range(MIN_K, MAX_K + 1) for c in itertools.combinations(ALL_BLOCKS, k)]
combo_rng = np.random.RandomState(seed=0)      # the seeded shuffle behind the train / test split of the combinations
combo_rng.shuffle(_all)
ALL_COMBINATIONS = _all
_cut = int(len(_all) * 0.9)
TRAIN_COMBINATIONS, TEST_COMBINATIONS = _all[:_cut], _all[_cut:]

RGB = {"red": (214, 39, 40), "blue": (31, 119, 180), "green": (44, 160, 44), "yellow": (255, 221, 51), "purple": (148, 103, 189)}


def _variant(mode):
    return mode if isinstance(mode, LanguageTableBlockVariants) else LanguageTableBlockVariants(mode)


def get_all_block_subsets(mode, training):
    mode = _variant(mode)
    V = LanguageTableBlockVariants
    table = {V.BLOCK_1: [FIXED_1_COMBINATION], V.BLOCK_4: [FIXED_4_COMBINATION], V.BLOCK_8: [FIXED_8_COMBINATION],
             V.BLOCK_4_WPOLE: [FIXED_4_COMBINATION_WPOLE], V.BLOCK_8_WPOLE: [FIXED_8_COMBINATION_WPOLE]}
    if mode == V.N_CHOOSE_K:
        return TRAIN_COMBINATIONS if training else TEST_COMBINATIONS
    return table[mode]


def get_block_set(mode):
    mode = _variant(mode)
    V = LanguageTableBlockVariants
    table = {V.BLOCK_1: FIXED_1_COMBINATION, V.BLOCK_4: FIXED_4_COMBINATION, V.BLOCK_8: FIXED_8_COMBINATION,
             V.N_CHOOSE_K: ALL_BLOCKS}
    if mode not in table:
        raise ValueError("Unsupported block mode")
    return table[mode]


def get_all_block_pairs(mode):
    return itertools.permutations(get_block_set(mode), 2)


def get_blocks_text_descriptions(mode):
    return [" ".join(b.split("_")) for b in get_block_set(mode)]
