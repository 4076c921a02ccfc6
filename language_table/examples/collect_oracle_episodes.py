"""Episodes of the scripted RRT* oracle solving Language-Table tasks, in the RLDS nesting of the published datasets.

Upstream, the ``language_table_*_oracle_sim`` datasets were produced by running the push oracle in the simulator and logging
every control step; training here normally starts from a converted copy of such a dataset (``rlds_np_convert.py --tfds_dir``).
Offline there is no copy to convert — this module makes the episodes instead: reset the environment until the oracle finds a
plan (what ``eval/main_rt1.py`` does before every roll-out), let it push until the reward reports success or the step limit
is hit, keep successful episodes only, and yield them as ``{"steps": [{"observation": {rgb, instruction (byte code),
effector_translation, effector_target_translation}, "action", "reward", "is_first", "is_last", "is_terminal"}, ...]}`` —
what ``rt1b200.data.episodes.convert_rlds`` (and the reference's converter) consumes.

    python rlds_np_convert.py --out data --oracle 200 --train 180 --val 10 --test 10        # collect + embed + split
    python -m language_table.examples.collect_oracle_episodes --out exported --episodes 20    # RLDS-nested .npy exports
"""
import argparse
import os
from typing import Dict, Iterator, Optional

import numpy as np

REWARDS = ("block2block", "block2absolutelocation", "block2relativelocation", "block2block_relative_location", "separate_blocks",
           "block1_to_corner")


def _reward_factory(name: str):
    import importlib
    mod = importlib.import_module("language_table.environments.rewards." + name)
    classes = {"block2block": "BlockToBlockReward", "block2absolutelocation": "BlockToAbsoluteLocationReward",
               "block2relativelocation": "BlockToRelativeLocationReward",
               "block2block_relative_location": "BlockToBlockRelativeLocationReward", "separate_blocks": "SeparateBlocksReward",
               "block1_to_corner": "Block1ToCornerLocationReward"}
    return getattr(mod, classes[name])


def oracle_episodes(num_episodes: int, reward: str = "block2block", block_mode: str = "BLOCK_8", seed: int = 0,
                    max_episode_steps: int = 200, action_noise_std: float = 0.0, keep_failures: bool = False,
                    progress: Optional[int] = None) -> Iterator[Dict]:
    """Yields ``num_episodes`` RLDS-nested episodes.  One environment and one oracle are reused; the board, the task and the
    oracle's random stream all derive from ``seed``, so the same call reproduces the same episodes."""
    from language_table.environments import blocks, language_table
    from language_table.environments.oracles import push_oracle_rrt_slowdown
    env = language_table.LanguageTable(block_mode=getattr(blocks.LanguageTableBlockVariants, block_mode),
                                       reward_factory=_reward_factory(reward), seed=seed, render_text_in_image=False)
    oracle = push_oracle_rrt_slowdown.ObstacleOrientedPushOracleBoard2dRRT(env, rng=np.random.RandomState(seed + 1),
                                                                          action_noise_std=action_noise_std)
    made = attempts = 0
    while made < num_episodes:
        attempts += 1
        if attempts > 20 * num_episodes + 100:
            raise RuntimeError(f"only {made} of {num_episodes} episodes after {attempts} attempts: is the task solvable?")
        for _ in range(20):                                   # a board the oracle can plan on
            obs = env.reset()
            oracle.reset()
            if oracle.get_plan(env._compute_state(render=False)):
                break
        else:
            continue
        steps, done, reward_value = [], False, 0.0
        for t in range(max_episode_steps):
            action = np.asarray(oracle.action(env._compute_state(render=False)), np.float32)
            steps.append({"observation": {k: np.asarray(obs[k]) for k in ("rgb", "instruction", "effector_translation",
                                                                          "effector_target_translation")},
                          "action": action, "reward": np.float32(reward_value), "is_first": t == 0, "is_last": False,
                          "is_terminal": False})
            obs, reward_value, done, _ = env.step(action)
            if done:
                break
        if not (done or keep_failures):
            continue
        # the terminal step carries the final observation and the reward that ended the episode (RLDS convention)
        steps.append({"observation": {k: np.asarray(obs[k]) for k in ("rgb", "instruction", "effector_translation",
                                                                      "effector_target_translation")},
                      "action": np.zeros(2, np.float32), "reward": np.float32(reward_value), "is_first": False, "is_last": True,
                      "is_terminal": bool(done)})
        made += 1
        if progress and made % progress == 0:
            print(f"[oracle] {made} / {num_episodes} episodes ({attempts} attempts)", flush=True)
        yield {"steps": steps}


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__.split("\n\n")[0])
    ap.add_argument("--out", required=True, help="directory for episode_<n>.npy exports (RLDS nesting; feed to rlds_np_convert.py --source)")
    ap.add_argument("--episodes", type=int, default=20)
    ap.add_argument("--reward", choices=REWARDS, default="block2block")
    ap.add_argument("--block_mode", default="BLOCK_8")
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--max_episode_steps", type=int, default=200)
    a = ap.parse_args(argv)
    os.makedirs(a.out, exist_ok=True)
    n = 0
    for n, ep in enumerate(oracle_episodes(a.episodes, a.reward, a.block_mode, a.seed, a.max_episode_steps, progress=10)):
        np.save(os.path.join(a.out, f"episode_{n}.npy"), np.array(ep, dtype=object), allow_pickle=True)
    print(f"wrote {n + 1} episodes to {a.out}")
    return n + 1


if __name__ == "__main__":
    main()
