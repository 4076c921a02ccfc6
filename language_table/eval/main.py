"""Closed-loop evaluation of a LAVA BC checkpoint on the five Language-Table task families (reference
``language_table/eval/main.py:46-160``: 50 episodes x 5 rewards x <= 200 steps, BLOCK_8, seed 0, initial boards re-sampled
until the RRT* oracle finds a plan).  Upstream this script is broken in the fork (it references removed wrappers); here it
runs on the repository's environment, wrappers and ``BCJaxPyPolicy``.

    python -m language_table.eval.main --config CONFIG.py --checkpoint_path CKPT.pt --workdir DIR [--num_evals_per_reward 50]
"""
import argparse
import collections
import json
import os

import numpy as np

from language_table.environments import blocks, language_table
from language_table.environments.oracles import push_oracle_rrt_slowdown
from language_table.environments.rewards import (block2absolutelocation, block2block, block2block_relative_location,
                                                 block2relativelocation, separate_blocks)
from language_table.eval import wrappers
from language_table.train import main as train_main
from language_table.train import train as train_lib
from language_table.train.policy import BCJaxPyPolicy

REWARDS = collections.OrderedDict([
    ("blocktoblock", block2block.BlockToBlockReward),
    ("blocktoabsolutelocation", block2absolutelocation.BlockToAbsoluteLocationReward),
    ("blocktoblockrelativelocation", block2block_relative_location.BlockToBlockRelativeLocationReward),
    ("blocktorelativelocation", block2relativelocation.BlockToRelativeLocationReward),
    ("separate", separate_blocks.SeparateBlocksReward),
])


def evaluate_checkpoint(checkpoint_path, workdir, config, num_evals_per_reward=50, max_episode_steps=200, device=None, seed=0,
                        save_frames=False):
    os.makedirs(workdir, exist_ok=True)
    model = train_lib.create_model(config)
    policy = BCJaxPyPolicy.from_checkpoint(model, checkpoint_path, device=device)
    results = collections.OrderedDict()
    for name, factory in REWARDS.items():
        base = language_table.LanguageTable(block_mode=blocks.LanguageTableBlockVariants.BLOCK_8, reward_factory=factory, seed=seed)
        # language: the stored sentence embedding, or BPE ids for the in-graph CLIP tower (upstream's ClipTokenWrapper)
        if config.model.get("lava_lang_encoder") == "clip":
            worded = wrappers.ClipTokenWrapper(base, bpe_path=config.get("clip_bpe_path"))
        else:
            worded = wrappers.UseTokenWrapper(base)
        env = wrappers.HistoryWrapper(
            wrappers.CentralCropImageWrapper(worded, config.data_target_height, config.data_target_width,
                                             config.random_crop_factor),
            history_length=config.sequence_length, tile_first_step_obs=True)
        oracle = push_oracle_rrt_slowdown.ObstacleOrientedPushOracleBoard2dRRT(base, rng=np.random.RandomState(seed))
        wins = 0
        for ep in range(num_evals_per_reward):
            for _ in range(20):
                obs = env.reset()
                oracle.reset()
                if oracle.get_plan(base._compute_state()):
                    break
            frames = [base.render()] if save_frames else []
            for _ in range(max_episode_steps):
                obs, _, done, _ = env.step(policy.action(obs))
                if save_frames:
                    frames.append(base.render())
                if done:
                    break
            wins += int(base.succeeded)
            if save_frames:
                tag = "success" if base.succeeded else "failure"
                np.save(os.path.join(workdir, f"{name}_{ep}_{tag}.npy"), np.stack(frames))
        results[name] = {"episodes": num_evals_per_reward, "successes": wins}
    print(json.dumps(results))
    with open(os.path.join(workdir, "results.json"), "w") as f:
        json.dump(results, f, indent=1)
    return results


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", required=True)
    ap.add_argument("--checkpoint_path", required=True)
    ap.add_argument("--workdir", required=True)
    ap.add_argument("--num_evals_per_reward", type=int, default=50)
    ap.add_argument("--max_episode_steps", type=int, default=200)
    ap.add_argument("--device", default=None)
    a, extra = ap.parse_known_args(argv)
    config = train_main.apply_config_overrides(train_main.load_config(a.config), extra)     # --config.<field>=<value>, as in training
    return evaluate_checkpoint(a.checkpoint_path, a.workdir, config, a.num_evals_per_reward, a.max_episode_steps, a.device)


if __name__ == "__main__":
    main()
