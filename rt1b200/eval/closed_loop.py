"""Batched closed-loop evaluation: E synthetic boards stepped against ONE batched RT-1 policy call per control step.

The reference evaluates one environment at a time (batch-1 inference, CPU simulation and two software renders per
step, ``language_table/eval/main_rt1.py:152-199``).  Here the boards, the rasteriser, the crop/resize and the policy all
live on the GPU: a control step is render → crop/resize → ``policy_step`` → clip → board step, with per-environment
rolling state and no host round trip except the optional latency read-out.  Per-step latency is measured with CUDA
events (p50 / p99 are BASELINE config 4).
"""
from __future__ import annotations

from typing import Dict, List, Optional

import numpy as np
import torch

from ..envs.board import BatchedLanguageTable


@torch.no_grad()
def run_closed_loop(model, num_envs: int = 64, episodes: int = 1, max_episode_steps: int = 80, device="cuda", seed: int = 0,
                    action_clip: float = 0.03, image_hw=(256, 456), record_frames: bool = False, use_oracle: bool = False,
                    block_mode=None, graphed: bool = False) -> Dict:
    """``graphed=True`` replays ONE captured CUDA graph per control step for the policy (``rt1b200.eval.graphed``; fixed
    centred crop instead of the reference's per-call random pad-shift)."""
    from language_table.environments import blocks as blocks_module
    dev = torch.device(device)
    env = BatchedLanguageTable(num_envs, block_mode or blocks_module.LanguageTableBlockVariants.BLOCK_8, device=dev, seed=seed,
                               max_episode_steps=max_episode_steps, image_hw=image_hw)
    dtype = next(model.parameters()).dtype
    model.eval()
    successes, lengths, latencies_ms, frames = 0, [], [], []
    timing = dev.type == "cuda"
    gp = None
    if graphed and not use_oracle and dev.type == "cuda":
        from .graphed import GraphedPolicy
        gp = GraphedPolicy(model, num_envs, image_hw[0], image_hw[1], device=dev)
    for ep in range(episodes):
        env.reset()
        state = model.initial_state(num_envs, device=dev, dtype=dtype)
        if gp is not None:
            gp.reset()
        alive = torch.ones(num_envs, dtype=torch.bool, device=dev)
        ep_len = torch.zeros(num_envs, dtype=torch.long, device=dev)
        for t in range(max_episode_steps):
            if timing:
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
            obs = env.observe()
            if use_oracle:
                action = env.scripted_action(action_clip)
            elif gp is not None:
                action = gp.step(obs["image"], obs["natural_language_embedding"])["action"].float().clamp(-action_clip, action_clip)
            else:
                out, state = model.policy_step({"image": obs["image"], "natural_language_embedding": obs["natural_language_embedding"]}, state)
                action = out["action"].float().clamp(-action_clip, action_clip)
            action = torch.where(alive[:, None], action, torch.zeros_like(action))
            _, done = env.step(action)
            if timing:
                e1.record()
                e1.synchronize()
                latencies_ms.append(e0.elapsed_time(e1))
            if record_frames:
                frames.append(obs["rgb"][0].cpu().numpy())
            ep_len += alive.long()
            alive &= ~done
            if not bool(alive.any()):
                break
        successes += int(env.succeeded.sum())
        lengths += ep_len.tolist()
    lat = np.asarray(latencies_ms[3:] or latencies_ms or [0.0])
    return {"episodes": episodes * num_envs, "successes": successes, "success_rate": successes / max(1, episodes * num_envs),
            "mean_episode_length": float(np.mean(lengths)) if lengths else 0.0,
            "step_latency_ms_p50": float(np.percentile(lat, 50)), "step_latency_ms_p99": float(np.percentile(lat, 99)),
            "num_envs": num_envs, "frames": frames}
