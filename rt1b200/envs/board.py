"""Batched synthetic Language-Table board: state, kinematic push dynamics, block-to-block reward, rendering.

Device-resident replacement for one ``LanguageTable`` PyBullet environment per process (reference
``language_table/environments/language_table.py``): E boards advance in lock-step as tensors, so the closed-loop
evaluator can run 64 environments against one batched policy step (BASELINE config 4).  Bullet's rigid-body contact
model is replaced by a documented stand-in: the effector and the blocks are discs; a 10 Hz control step moves the
effector to its (workspace-clipped) target in sub-steps and resolves overlaps by pushing blocks out along the contact
normal, block-block overlaps symmetrically.  Constants, reset rejection sampling, the 0.05 m block-to-block success
radius and the 80-step episode cap follow the reference (``constants.py``, ``language_table.py:822-931``,
``rewards/block2block.py``, ``eval/main_rt1.py:118-119``).
"""
from __future__ import annotations

from typing import Dict, Optional, Sequence

import torch

from language_table.environments import blocks as blocks_module
from language_table.environments import constants as C

from . import raster
from .scene import SceneBuilder

TARGET_BLOCK_DISTANCE = 0.05
SUBSTEPS = 4


def resolve_contacts(eff: torch.Tensor, blocks: torch.Tensor, mask: torch.Tensor, iters: int = 3):
    """Push blocks out of the effector disc and out of each other. eff [E,2], blocks [E,B,2], mask [E,B]."""
    r_eb = C.EFFECTOR_RADIUS + C.BLOCK_RADIUS
    r_bb = 2 * C.BLOCK_RADIUS
    for _ in range(iters):
        d = blocks - eff[:, None, :]
        dist = d.norm(dim=-1, keepdim=True).clamp_min(1e-6)
        push = (r_eb - dist).clamp_min(0.0) * (d / dist)
        blocks = blocks + push * mask[..., None]
        diff = blocks[:, :, None, :] - blocks[:, None, :, :]                      # [E,B,B,2]
        dd = diff.norm(dim=-1, keepdim=True).clamp_min(1e-6)
        both = (mask[:, :, None] & mask[:, None, :])[..., None]
        eye = torch.eye(blocks.shape[1], dtype=torch.bool, device=blocks.device)[None, :, :, None]
        overlap = ((r_bb - dd).clamp_min(0.0) * (diff / dd)) * (both & ~eye)
        blocks = blocks + 0.5 * overlap.sum(dim=2)
    lo = torch.tensor([C.X_MIN - 0.02, C.Y_MIN - 0.02], device=blocks.device)
    hi = torch.tensor([C.X_MAX + 0.02, C.Y_MAX + 0.02], device=blocks.device)
    return torch.minimum(torch.maximum(blocks, lo), hi)


class BatchedLanguageTable:
    def __init__(self, num_envs: int, block_mode=blocks_module.LanguageTableBlockVariants.BLOCK_8, device="cpu", seed: int = 0,
                 max_episode_steps: int = 80, image_hw=(256, 456), crop: float = 0.95, add_invisible_walls: bool = False,
                 reset_buffer: float = C.WORKSPACE_BOUNDS_BUFFER):
        self.E = num_envs
        # initial effector and block positions are drawn this far inside the workspace (reference language_table.py:837-840)
        self.reset_buffer = float(reset_buffer)
        # The reference's experimental "invisible walls" (language_table.py:933-990): four static boxes of half-width 0.04 m
        # just outside the workspace that keep blocks from sliding off the board.  Here: block centres are clamped to the
        # rectangle the walls' inner faces leave, shrunk by the block radius.
        self.add_invisible_walls = bool(add_invisible_walls)
        self.device = torch.device(device)
        self.block_names = list(blocks_module.get_block_set(block_mode))
        self.B = len(self.block_names)
        self.gen = torch.Generator(device="cpu").manual_seed(seed)
        self.max_episode_steps = max_episode_steps
        self.image_hw, self.crop = image_hw, crop
        self.scene = SceneBuilder(self.block_names, self.device)
        z = lambda *s, dt=torch.float32: torch.zeros(*s, dtype=dt, device=self.device)   # noqa: E731
        self.effector, self.effector_target = z(num_envs, 2), z(num_envs, 2)
        self.block_xy, self.block_yaw = z(num_envs, self.B, 2), z(num_envs, self.B)
        self.block_mask = torch.ones(num_envs, self.B, dtype=torch.bool, device=self.device)
        self.start_block = z(num_envs, dt=torch.long)
        self.target_block = z(num_envs, dt=torch.long)
        self.instruction_embedding = z(num_envs, 512)
        self.steps = z(num_envs, dt=torch.long)
        self.succeeded = torch.zeros(num_envs, dtype=torch.bool, device=self.device)
        self.reset()

    # ------------------------------------------------------------------------------------------------ reset
    def _rand(self, *shape):
        return torch.rand(*shape, generator=self.gen).to(self.device)

    def reset(self, env_ids: Optional[torch.Tensor] = None):
        """Rejection-sample block poses (no two blocks closer than a diameter + margin, none under the effector) and a
        block-to-block task whose blocks start at least 0.06 m apart; fresh random instruction embedding."""
        ids = torch.arange(self.E, device=self.device) if env_ids is None else env_ids.to(self.device)
        n = len(ids)
        if n == 0:
            return
        lo = torch.tensor([C.X_MIN + self.reset_buffer, C.Y_MIN + self.reset_buffer], device=self.device)
        hi = torch.tensor([C.X_MAX - self.reset_buffer, C.Y_MAX - self.reset_buffer], device=self.device)
        eff = lo + self._rand(n, 2) * (hi - lo)
        xy = lo + self._rand(n, self.B, 2) * (hi - lo)
        for _ in range(30):                                    # resample offenders until every board is valid
            d = (xy[:, :, None] - xy[:, None, :]).norm(dim=-1) + torch.eye(self.B, device=self.device) * 10
            bad = (d.min(dim=-1).values < 2 * C.BLOCK_RADIUS + 0.015) | ((xy - eff[:, None]).norm(dim=-1) < C.ARM_DISTANCE_THRESHOLD)
            if not bool(bad.any()):
                break
            xy = torch.where(bad[..., None], lo + self._rand(n, self.B, 2) * (hi - lo), xy)
        xy = resolve_contacts(eff, xy, torch.ones(n, self.B, dtype=torch.bool, device=self.device), iters=4)
        start = torch.randint(0, self.B, (n,), generator=self.gen).to(self.device)
        offs = torch.randint(1, self.B, (n,), generator=self.gen).to(self.device) if self.B > 1 else torch.zeros_like(start)
        target = (start + offs) % self.B
        self.effector[ids], self.effector_target[ids] = eff, eff.clone()
        self.block_xy[ids], self.block_yaw[ids] = xy, self._rand(n, self.B) * 6.2832
        self.start_block[ids], self.target_block[ids] = start, target
        emb = torch.randn(n, 512, generator=self.gen).to(self.device)
        self.instruction_embedding[ids] = emb / emb.norm(dim=-1, keepdim=True)
        self.steps[ids] = 0
        self.succeeded[ids] = False

    # ------------------------------------------------------------------------------------------------- step
    def task_distance(self) -> torch.Tensor:
        idx = torch.arange(self.E, device=self.device)
        return (self.block_xy[idx, self.start_block] - self.block_xy[idx, self.target_block]).norm(dim=-1)

    def step(self, action: torch.Tensor, effector_path: Optional[torch.Tensor] = None):
        """``action`` [E,2] metres (Δx, Δy) added to the effector target. Returns (reward [E], done [E]).
        ``effector_path`` [E, n, 2] (optional): where the effector actually is at n instants of this control step — an arm
        model that lags its target supplies it (``LanguageTable`` with ``arm_kinematics``); without it the effector
        reaches the target by linear interpolation."""
        action = action.to(self.device, torch.float32)
        lo = torch.tensor([C.X_MIN, C.Y_MIN], device=self.device)
        hi = torch.tensor([C.X_MAX, C.Y_MAX], device=self.device)
        self.effector_target = torch.minimum(torch.maximum(self.effector_target + action, lo), hi)
        if effector_path is None:
            frac = torch.arange(1, SUBSTEPS + 1, device=self.device, dtype=torch.float32)[None, :, None] / SUBSTEPS
            effector_path = self.effector[:, None] + (self.effector_target - self.effector)[:, None] * frac
        for k in range(effector_path.shape[1]):
            self.effector = effector_path[:, k].to(self.device, torch.float32)
            self.block_xy = resolve_contacts(self.effector, self.block_xy, self.block_mask, iters=2)
            if self.add_invisible_walls:
                r = C.BLOCK_RADIUS
                wlo = torch.tensor([C.X_MIN + 0.02 + r, C.Y_MIN + 0.03 + r], device=self.device)
                whi = torch.tensor([C.X_MAX - 0.05 - r, C.Y_MAX - 0.03 - r], device=self.device)
                self.block_xy = torch.minimum(torch.maximum(self.block_xy, wlo), whi)
        self.steps += 1
        success = self.task_distance() < TARGET_BLOCK_DISTANCE
        self.succeeded |= success
        reward = success.float() * 100.0
        done = success | (self.steps >= self.max_episode_steps)
        return reward, done

    # --------------------------------------------------------------------------------------------- observe
    def render(self) -> torch.Tensor:
        """uint8 [E,180,320,3] frames in a device buffer (no window, no graphics API)."""
        tris, ntris = self.scene.triangles(self.effector, self.block_xy, self.block_yaw, self.block_mask)
        return raster.render(tris, ntris)

    def observe(self) -> Dict[str, torch.Tensor]:
        """Policy inputs: ``image`` float32 [E,3,H,W] in [0,1] (centre crop + resize) and the instruction embedding."""
        frames = self.render()
        return {"image": raster.crop_resize(frames, self.image_hw[0], self.image_hw[1], self.crop),
                "natural_language_embedding": self.instruction_embedding, "rgb": frames}

    def scripted_action(self, max_step: float = 0.03) -> torch.Tensor:
        """A simple oracle: go behind the start block (opposite the target), then push through it."""
        idx = torch.arange(self.E, device=self.device)
        s, t = self.block_xy[idx, self.start_block], self.block_xy[idx, self.target_block]
        to_t = t - s
        to_t = to_t / to_t.norm(dim=-1, keepdim=True).clamp_min(1e-6)
        behind = s - to_t * (C.BLOCK_RADIUS + C.EFFECTOR_RADIUS + 0.015)
        off = self.effector - behind
        aligned = (off.norm(dim=-1) < 0.012) | (((self.effector - s) * to_t).sum(-1) < -0.5 * (C.BLOCK_RADIUS + C.EFFECTOR_RADIUS))
        lateral = (self.effector - s) - ((self.effector - s) * to_t).sum(-1, keepdim=True) * to_t
        aligned &= lateral.norm(dim=-1) < 0.012
        goal = torch.where(aligned[:, None], s + to_t * 0.02, behind)
        # when approaching from the front, swing around the block instead of pushing it away
        in_front = ((self.effector - s) * to_t).sum(-1) > -0.01
        side = torch.stack([-to_t[:, 1], to_t[:, 0]], -1)
        sign = torch.sign((lateral * side).sum(-1, keepdim=True))
        sign = torch.where(sign == 0, torch.ones_like(sign), sign)
        detour = s + side * sign * 0.07 - to_t * 0.02
        goal = torch.where((in_front & ~aligned & (lateral.norm(dim=-1) < 0.06))[:, None], detour, goal)
        step = goal - self.effector_target
        n = step.norm(dim=-1, keepdim=True).clamp_min(1e-6)
        return step / n * torch.minimum(n, torch.full_like(n, max_step))
