"""RT-1 policy adapter for the closed-loop evaluator (reference ``language_table/train/policy.py:32``,
``BCJaxPyPolicyRT1``): last frame of the history stack + sentence embedding → ``TransformerNetwork`` inference call with
a rolling network state → de-normalise → clip to ±0.03 m."""
from typing import Dict, Optional

import numpy as np
import torch

EPS = np.finfo(np.float32).eps


class BCJaxPyPolicyRT1:
    def __init__(self, model, network_state: Optional[Dict[str, torch.Tensor]] = None, device=None, action_std=1.0,
                 action_mean=0.0, action_clip=0.03):
        self.model = model
        self.device = torch.device(device) if device is not None else next(model.parameters()).device
        self.network_state = network_state
        self.action_mean, self.action_std = action_mean, max(float(action_std), EPS)
        self.action_clip = action_clip

    def reset(self, batch_size: int = 1):
        dtype = next(self.model.parameters()).dtype
        self.network_state = self.model.initial_state(batch_size, device=self.device, dtype=dtype)

    def action(self, observation) -> np.ndarray:
        """``observation``: history-stacked dict with ``rgb_sequence [T,H,W,3]`` and ``instruction_tokenized_use [T,512]``."""
        if self.network_state is None:
            self.reset(1)
        rgb = np.asarray(observation["rgb_sequence"])[-1]
        emb = np.asarray(observation["instruction_tokenized_use"])[-1]
        obs = {"image": torch.from_numpy(np.ascontiguousarray(rgb)).permute(2, 0, 1)[None].to(self.device, torch.float32),
               "natural_language_embedding": torch.from_numpy(emb)[None].to(self.device, torch.float32)}
        with torch.no_grad():
            out, self.network_state = self.model.policy_step(obs, self.network_state)
        act = out["action"][0].float().cpu().numpy() * self.action_std + self.action_mean
        return np.clip(act, -self.action_clip, self.action_clip).astype(np.float32)


class BCJaxPyPolicy:
    """Policy adapter for the LAVA BC baseline (the class the reference's ``eval/main.py:95`` expects; upstream it wraps a
    Flax checkpoint, here a ``language_table.train.bc.BCAgent`` state): stacked frames + sentence embedding → model →
    de-normalise with the stored action statistics → clip to the dataset's action range."""

    def __init__(self, model, norm_info: Dict, device=None, action_clip: Optional[float] = None):
        self.model = model.eval()
        self.device = torch.device(device) if device is not None else next(model.parameters()).device
        st = norm_info["action_statistics"]
        self.mean, self.std = np.asarray(st["mean"], np.float32), np.maximum(np.asarray(st["std"], np.float32), EPS)
        self.amin, self.amax = np.asarray(norm_info["action_min"], np.float32), np.asarray(norm_info["action_max"], np.float32)
        self.action_clip = action_clip

    @classmethod
    def from_checkpoint(cls, model, checkpoint_path: str, device=None, **kw):
        sd = torch.load(checkpoint_path, map_location=device or "cpu", weights_only=False)
        model.load_state_dict(sd["params"])
        return cls(model.to(device) if device is not None else model, sd["norm_info"], device=device, **kw)

    def action(self, observation) -> np.ndarray:
        """``observation``: history-stacked dict with ``rgb_sequence [T,H,W,3]`` in [0,1] and either ``instruction_tokenized_use
        [T,512]`` (sentence embedding) or ``instruction_tokenized_clip [T,77]`` (BPE ids for a model with the in-graph CLIP tower)."""
        obs = {"rgb": torch.from_numpy(np.ascontiguousarray(observation["rgb_sequence"]))[None].to(self.device, torch.float32)}
        if "instruction_tokenized_clip" in observation:
            obs["instruction_tokenized_clip"] = torch.from_numpy(np.asarray(observation["instruction_tokenized_clip"], np.int64))[None].to(self.device)
        else:
            obs["clip_embedding"] = torch.from_numpy(np.asarray(observation["instruction_tokenized_use"], np.float32))[None].to(self.device)
        with torch.no_grad():
            a = self.model(obs)[0].float().cpu().numpy()
        a = np.clip(a * self.std + self.mean, self.amin, self.amax)
        if self.action_clip is not None:
            a = np.clip(a, -self.action_clip, self.action_clip)
        return a.astype(np.float32)
