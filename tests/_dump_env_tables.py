"""Dumps the deterministic tables of a Language-Table package (whichever ``language_table`` is first on sys.path) as JSON:
every instruction string per block mode, block sets and their seeded train / test split, synonym tables, named board
locations, numeric constants.  tests/test_reference_parity_data.py runs it once against the reference and once against this
repository and compares the two files."""
import json
import sys

import numpy as np


def plain(v):
    if isinstance(v, np.ndarray):
        return [plain(x) for x in v.tolist()]
    if isinstance(v, (np.floating, np.integer)):
        return v.item()
    if isinstance(v, dict):
        return {str(k): plain(x) for k, x in v.items()}
    if isinstance(v, (list, tuple)):
        return [plain(x) for x in v]
    if isinstance(v, float):
        return round(v, 9)
    return v


def numeric_constants(mod):
    out = {}
    for k in dir(mod):
        if not k.isupper():
            continue
        v = getattr(mod, k)
        if isinstance(v, (int, float, str, bool, list, tuple, dict, np.ndarray)) and not k.endswith("_PATH") and not k.endswith("_PATHS"):
            try:
                out[k] = plain(v)
                json.dumps(out[k])
            except (TypeError, ValueError):
                out.pop(k, None)
    return out


def main(path):
    from language_table.environments import blocks, constants
    from language_table.environments.rewards import (block2absolutelocation, block2block_relative_location, block2relativelocation,
                                                     constants as reward_constants, instructions, synonyms)
    modes = blocks.LanguageTableBlockVariants
    out = {"instructions": {m: sorted(instructions.generate_all_instructions(getattr(modes, m)))
                            for m in ("BLOCK_4", "BLOCK_8", "N_CHOOSE_K")},
           "blocks": {"ALL_BLOCKS": list(blocks.ALL_BLOCKS), "FIXED_4": list(blocks.FIXED_4_COMBINATION),
                      "FIXED_8": list(blocks.FIXED_8_COMBINATION),
                      "TRAIN": [list(c) for c in blocks.TRAIN_COMBINATIONS], "TEST": [list(c) for c in blocks.TEST_COMBINATIONS],
                      "text": {m: list(blocks.get_blocks_text_descriptions(getattr(modes, m))) for m in ("BLOCK_4", "BLOCK_8", "N_CHOOSE_K")}},
           "constants": numeric_constants(constants), "reward_constants": numeric_constants(reward_constants),
           "absolute_locations": plain(block2absolutelocation.ABSOLUTE_LOCATIONS),
           "location_synonyms": plain(block2absolutelocation.LOCATION_SYNONYMS),
           "relative_directions": plain({k: np.asarray(v, float) for k, v in block2relativelocation.DIRECTIONS.items()}),
           "b2b_relative": {"DIRECTIONS": plain(block2block_relative_location.DIRECTIONS),
                            "tasks": plain(block2block_relative_location.UNIQUE_TASK_STRINGS)},
           "synonyms": {k: plain(getattr(synonyms, k)) for k in dir(synonyms) if k.isupper()}}
    # the small functions around the tables
    from language_table.environments.rewards import reward as reward_base
    sets = {"fixed8": list(blocks.FIXED_8_COMBINATION), "fixed4": list(blocks.FIXED_4_COMBINATION),
            "subset": list(blocks.TRAIN_COMBINATIONS[3])}
    grid = [(round(x, 3), round(y, 3)) for x in np.linspace(0.1, 0.65, 12) for y in np.linspace(-0.35, 0.35, 15)]
    rel_rng = np.random.RandomState(0)
    pairs = [(rel_rng.uniform(0.2, 0.5, 2), None) for _ in range(200)]
    pairs = [(a, a + rel_rng.uniform(-0.12, 0.12, 2)) for a, _ in pairs]
    out["functions"] = {
        "block_synonyms": {k: {b: plain(synonyms.get_block_synonyms(b, v)) for b in v} for k, v in sets.items()},
        "count_color_shape": {k: plain([dict(c) for c in synonyms.count_color_shape(v)]) for k, v in sets.items()},
        "color_shape": {b: plain(synonyms.get_color_shape(b)) for b in blocks.ALL_BLOCKS},
        "block_set": {m: plain(sorted(blocks.get_block_set(getattr(modes, m)))) for m in ("BLOCK_4", "BLOCK_8", "N_CHOOSE_K")},
        "block_pairs": {m: plain([list(p) for p in blocks.get_all_block_pairs(getattr(modes, m))]) for m in ("BLOCK_4", "BLOCK_8")},
        "subsets": {"%s_%s" % (m, t): len(blocks.get_all_block_subsets(getattr(modes, m), t))
                    for m in ("BLOCK_4", "BLOCK_8", "N_CHOOSE_K") for t in (True, False)},
        "inside_bounds": [bool(reward_base.target_inside_bounds(np.array(p))) for p in grid],
        "is_b2b_relative_pair": [bool(block2block_relative_location.is_block2block_relative_pair(a, b)) for a, b in pairs]}
    from language_table.environments.utils.pose3d import Pose3d
    from scipy.spatial import transform
    pose = Pose3d(rotation=transform.Rotation.from_euler("xyz", [0.1, -0.2, 0.3]), translation=np.array([1.0, 2.0, 3.0]))
    out["pose3d"] = {"vec7": plain(np.round(pose.vec7, 9)), "serialize": plain(pose.serialize()),
                     "roundtrip": plain(np.round(Pose3d.deserialize(pose.serialize()).vec7, 9)),
                     "asdict_keys": sorted(pose.asdict())}
    try:
        from language_table.environments.rewards import play
        rng = np.random.RandomState(0)
        out["play"] = {"train4": play.get_100_4block_instructions(), "test4": play.get_100_4block_instructions(return_train=False),
                       "families8": {fn: sorted(getattr(play, fn)()) for fn in (
                           "get_sort_tasks", "group_color_pairs", "get_colors_in_lines", "group_color_pairs_in_locations",
                           "get_line_tasks", "get_surround_tasks", "all_blocks_in_location",
                           "k_blocks_in_location_i_rest_in_location_j", "get_shape_instructions")},
                       "colors_in_locations_head": play.colors_in_locations()[:50], "edge_head": play.blocks_in_order_outer_edge()[:50]}
    except ImportError as e:                                   # the reference's play.py needs gin
        out["play"] = {"unavailable": str(e)}
    with open(path, "w") as f:
        json.dump(out, f)


if __name__ == "__main__":
    main(sys.argv[1])
