"""Closed-loop evaluation of an RT-1 checkpoint on the Language-Table board.

    python -m language_table.eval.main_rt1 --workdir=/tmp/rt1_eval [--ckpt PATH] [--num_envs 64] [--episodes 10]

Same entry point as the reference (``language_table/eval/main_rt1.py``: ``--workdir``, BLOCK_8 + block-to-block reward,
seed 0, episodes of at most 80 steps, actions clipped to ±0.03, success counted per episode, result dict printed).
Differences: the checkpoint path is a flag (upstream hard-codes it, ``:220``); ``--num_envs N > 1`` runs N boards on the
GPU against one batched policy call per step (upstream: one board, batch 1); every episode is written under ``--workdir``
as an animated PNG (``--video_format gif`` for GIF; upstream writes mp4 through imageio/ffmpeg, ``:196-199``, which are
not installed — the writers in ``rt1b200/utils/video.py`` need no codec).  Without ``--ckpt`` a randomly initialised policy is used,
which exercises the full loop (no trained checkpoint can be downloaded here).
"""
from __future__ import annotations

import argparse
import json
import os
from argparse import Namespace

import numpy as np
import torch


def get_ckpt_model(ckpt, device, dtype=None, height=256, width=456, seq_len=6):
    from rt1b200.train.trainer import RT1Module, load_policy_from_checkpoint
    args = Namespace(height=height, width=width, seq_len=seq_len)
    if ckpt:
        return load_policy_from_checkpoint(ckpt, args, device=device, dtype=dtype)
    module = RT1Module(args).to(device)
    if dtype is not None:
        from rt1b200.parallel.engine import keep_bn_stats_fp32
        module.to(dtype)
        keep_bn_stats_fp32(module)
    return module.eval()


# the reward families of the reference's table (``main_rt1.py:105-116``: "blocktoblock" enabled, the next four commented out),
# plus the single-block debugging task
REWARDS = {"blocktoblock": ("block2block", "BlockToBlockReward"),
           "blocktoabsolutelocation": ("block2absolutelocation", "BlockToAbsoluteLocationReward"),
           "blocktoblockrelativelocation": ("block2block_relative_location", "BlockToBlockRelativeLocationReward"),
           "blocktorelativelocation": ("block2relativelocation", "BlockToRelativeLocationReward"),
           "separate": ("separate_blocks", "SeparateBlocksReward"),
           "block1tocorner": ("block1_to_corner", "Block1ToCornerLocationReward")}


def evaluate_checkpoint(workdir, ckpt=None, num_envs=1, episodes=10, max_episode_steps=80, device=None, seed=0,
                        height=256, width=456, seq_len=6, use_oracle=False, video_format="png", rewards=("blocktoblock",),
                        block_mode="BLOCK_8"):
    from rt1b200.utils.video import write_video
    device = device or ("cuda" if torch.cuda.is_available() else "cpu")
    dtype = torch.bfloat16 if device.startswith("cuda") else None
    module = get_ckpt_model(ckpt, device, dtype, height, width, seq_len)
    module.model._eval_center_crop = True          # deterministic evaluation (upstream's eval crop is random)
    os.makedirs(workdir, exist_ok=True)
    results = {}
    rewards = tuple(rewards)
    unknown = [r for r in rewards if r not in REWARDS]
    if unknown:
        raise ValueError(f"unknown reward families {unknown}; known: {sorted(REWARDS)}")
    batched = num_envs > 1 or (device.startswith("cuda") and rewards == ("blocktoblock",) and block_mode == "BLOCK_8")
    if batched and (rewards != ("blocktoblock",) or block_mode != "BLOCK_8"):
        raise ValueError("the batched GPU boards implement the block-to-block task on BLOCK_8; use --num_envs 1 for other families")
    if batched:
        from rt1b200.eval.closed_loop import run_closed_loop
        n_rounds = max(1, (episodes + num_envs - 1) // num_envs)
        res = run_closed_loop(module.model, num_envs=num_envs, episodes=n_rounds, max_episode_steps=max_episode_steps, device=device,
                              seed=seed, image_hw=(height, width), record_frames=True, use_oracle=use_oracle)
        frames = res.pop("frames")
        if frames:
            write_video(os.path.join(workdir, f"blocktoblock_env0.{video_format}"), np.stack(frames), fps=10.0)
        results["blocktoblock"] = res
    else:
        # reference-shaped single-environment loop through the wrapper stack and the policy adapter
        import importlib
        from language_table.environments import blocks, language_table
        from language_table.environments.oracles import push_oracle_rrt_slowdown
        from language_table.eval import wrappers
        from language_table.train.policy import BCJaxPyPolicyRT1
        policy = BCJaxPyPolicyRT1(module.model, device=device)
        for name in rewards:
            mod, cls = REWARDS[name]
            factory = getattr(importlib.import_module("language_table.environments.rewards." + mod), cls)
            env = language_table.LanguageTable(block_mode=getattr(blocks.LanguageTableBlockVariants, block_mode),
                                               reward_factory=factory, seed=seed)
            base = env
            env = wrappers.HistoryWrapper(wrappers.CentralCropImageWrapper(wrappers.UseTokenWrapper(env), height, width, 0.95),
                                          history_length=seq_len, tile_first_step_obs=True)
            oracle = push_oracle_rrt_slowdown.ObstacleOrientedPushOracleBoard2dRRT(base, rng=np.random.RandomState(seed))
            successes = 0
            for ep in range(episodes):
                policy.reset(1)
                for _ in range(20):                   # re-sample the board until the oracle can plan it (upstream :162-172)
                    obs = env.reset()
                    oracle.reset()
                    if oracle.get_plan(base._compute_state(render=False)):
                        break
                frames = [base.render()]
                for t in range(max_episode_steps):
                    action = np.asarray(oracle.action(base._compute_state(render=False)), np.float32) if use_oracle else policy.action(obs)
                    obs, reward, done, _ = env.step(action)
                    frames.append(base.render())
                    if done:
                        break
                successes += int(base.succeeded)
                write_video(os.path.join(workdir, f"{name}_{ep}.{video_format}"), np.stack(frames), fps=10.0)
            results[name] = {"episodes": episodes, "successes": successes, "success_rate": successes / max(1, episodes)}
    print(json.dumps(results))
    with open(os.path.join(workdir, "results.json"), "w") as f:
        json.dump(results, f, indent=1)
    return results


def main(argv=None):
    p = argparse.ArgumentParser()
    p.add_argument("--workdir", type=str, required=True)
    p.add_argument("--ckpt", type=str, default=None)
    p.add_argument("--num_envs", type=int, default=1)
    p.add_argument("--episodes", type=int, default=10)
    p.add_argument("--max_episode_steps", type=int, default=80)
    p.add_argument("--device", type=str, default=None)
    p.add_argument("--seed", type=int, default=0)
    p.add_argument("--height", type=int, default=256)
    p.add_argument("--width", type=int, default=456)
    p.add_argument("--seq_len", type=int, default=6)
    p.add_argument("--oracle", action="store_true", help="drive the boards with the scripted oracle instead of the policy")
    p.add_argument("--video_format", choices=["png", "gif"], default="png", help="episode videos: animated PNG (lossless) or GIF")
    p.add_argument("--rewards", default="blocktoblock",
                   help="comma-separated task families (upstream's table: only blocktoblock enabled); known: " + ", ".join(sorted(REWARDS)))
    p.add_argument("--block_mode", default="BLOCK_8", help="BLOCK_1 / BLOCK_4 / BLOCK_8 / N_CHOOSE_K (single-environment path)")
    a = p.parse_args(argv)
    return evaluate_checkpoint(a.workdir, a.ckpt, a.num_envs, a.episodes, a.max_episode_steps, a.device, a.seed, a.height, a.width,
                               a.seq_len, a.oracle, a.video_format, [r for r in a.rewards.split(",") if r], a.block_mode)


if __name__ == "__main__":
    main()
