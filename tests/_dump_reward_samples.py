"""Seed-for-seed record of the reward classes of a Language-Table package (whichever ``language_table`` is first on sys.path):
for every pushing family and 8 seeds, the task ``reset`` samples on a fixed random board (instruction, blocks, target point) and
the ``(reward, done)`` it pays on five other boards.  tests/test_reference_parity_data.py compares the reference's record with
this repository's."""
import json, sys
import numpy as np
from language_table.environments import blocks
from language_table.environments.rewards import (block2block, block2absolutelocation, block2relativelocation,
    block2block_relative_location, separate_blocks, point2block, block1_to_corner)
modes = blocks.LanguageTableBlockVariants
def make_state(rng, names):
    st = {}
    for n in blocks.ALL_BLOCKS:
        st["block_%s_translation" % n] = np.array([rng.uniform(0.2, 0.55), rng.uniform(-0.25, 0.25)], np.float32)
        st["block_%s_orientation" % n] = np.array([rng.uniform(-3, 3)], np.float32)
    st["effector_translation"] = np.array([0.3, 0.0], np.float32)
    st["effector_target_translation"] = np.array([0.3, 0.0], np.float32)
    return st
out = {}
for name, mod, cls in [("b2b", block2block, "BlockToBlockReward"), ("b2abs", block2absolutelocation, "BlockToAbsoluteLocationReward"),
                       ("b2rel", block2relativelocation, "BlockToRelativeLocationReward"),
                       ("b2brel", block2block_relative_location, "BlockToBlockRelativeLocationReward"),
                       ("sep", separate_blocks, "SeparateBlocksReward"), ("p2b", point2block, "PointToBlockReward"),
                       ("corner", block1_to_corner, "Block1ToCornerLocationReward")]:
    rows = []
    for seed in range(8):
        srng = np.random.RandomState(1000 + seed)
        names = list(blocks.FIXED_8_COMBINATION)
        st = make_state(srng, names)
        r = getattr(mod, cls)(goal_reward=100.0, rng=np.random.RandomState(seed), delay_reward_steps=0, block_mode=modes.BLOCK_8)
        info = r.reset(st, names)
        row = {"info": "FAILURE" if info == "failure" else {k: (np.asarray(v).round(5).tolist() if not isinstance(v, (str, list)) else v) for k, v in vars(info).items()}}
        if info != "failure":
            rw = []
            for k in range(5):
                st2 = make_state(np.random.RandomState(2000 + seed * 10 + k), names)
                out_r = r.reward(st2)
                rw.append([float(out_r[0]), bool(out_r[1])])
            row["rewards"] = rw
            # and a board on which the task is done: the pushed block on its target (next to the target block; the effector on
            # the block for point-to-block), asked twice — the second call is what a delayed reward would pay
            solved = make_state(np.random.RandomState(3000 + seed), names)
            d = vars(info)
            start = d.get("block1") or d.get("block") or d.get("block_target")
            if d.get("target_translation") is not None:
                solved["block_%s_translation" % start] = np.asarray(d["target_translation"], np.float32)
            elif d.get("block2"):
                solved["block_%s_translation" % start] = solved["block_%s_translation" % d["block2"]] + np.array([0.03, 0.0], np.float32)
            else:
                solved["effector_translation"] = solved["block_%s_translation" % start].copy()
                solved["effector_target_translation"] = solved["block_%s_translation" % start].copy()
            row["solved"] = [[float(x[0]), bool(x[1])] for x in (r.reward(solved), r.reward(solved))]
        rows.append(row)
    out[name] = rows
json.dump(out, open(sys.argv[1], "w"))
