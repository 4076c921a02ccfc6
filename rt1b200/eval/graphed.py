"""CUDA-graphed closed-loop policy step.

A control step of the policy is ~600 kernel launches; at small batch (one robot = batch 1) the step is bound by the host
issuing them, not by the GPU.  ``GraphedPolicy`` captures ONE ``TransformerNetwork.policy_step`` — frame pre-processing,
FiLM-EfficientNet tokenizer, TokenLearner, the single decoder pass, arg-max, de-tokenisation and the roll of the per-environment
state — into a CUDA graph over static input/state buffers and replays it per control step: one launch from the host.

The reference applies a freshly sampled pad-shift crop on every call, also in evaluation
(``film_efficientnet/preprocessors.py:37-56`` via ``transformer_network.py:445``); a graph bakes its kernel arguments, so
the graphed policy uses a fixed offset: the centred window (``eval_center_crop``) unless ``crop_shift`` is given.
"""
from typing import Dict, Optional, Tuple

import torch


class GraphedPolicy:
    def __init__(self, model, batch_size: int, height: int, width: int, device=None, image_dtype=torch.float32,
                 crop_shift: Optional[Tuple[int, int]] = None, warmup: int = 3):
        self.model = model.eval()
        self.device = torch.device(device) if device is not None else next(model.parameters()).device
        assert self.device.type == "cuda", "CUDA graphs need a CUDA device"
        self.dtype = next(model.parameters()).dtype
        from ..models import preprocessors
        self._shift = tuple(crop_shift) if crop_shift is not None else preprocessors.crop_pads(height, width)
        self.obs = {"image": torch.zeros(batch_size, 3, height, width, dtype=image_dtype, device=self.device),
                    "natural_language_embedding": torch.zeros(batch_size, 512, dtype=torch.float32, device=self.device)}
        self.state = model.initial_state(batch_size, device=self.device, dtype=self.dtype)
        self.actions: Dict[str, torch.Tensor] = {}
        self._capture(warmup)

    def _one_step(self):
        prev = self.model._crop_shift_override
        self.model._crop_shift_override = self._shift
        try:
            act, new_state = self.model.policy_step(self.obs, self.state)
        finally:
            self.model._crop_shift_override = prev
        for k, v in new_state.items():                  # the roll of the rolling state is part of the captured work
            self.state[k].copy_(v.reshape(self.state[k].shape))
        return act

    def _capture(self, warmup: int):
        side = torch.cuda.Stream(device=self.device)
        side.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(side), torch.no_grad():
            for _ in range(max(1, warmup)):             # lazy one-time work (function attributes, masks, contexts) happens here
                self._one_step()
        torch.cuda.current_stream(self.device).wait_stream(side)
        torch.cuda.synchronize(self.device)
        self.reset()
        self.graph = torch.cuda.CUDAGraph()
        with torch.no_grad(), torch.cuda.graph(self.graph):
            act = self._one_step()
        self.actions = act
        self.reset()

    def reset(self, env_mask: Optional[torch.Tensor] = None):
        """Zero the rolling state of all environments, or of those selected by a bool mask ``[b]``."""
        for k, v in self.state.items():
            if env_mask is None:
                v.zero_()
            else:
                v[env_mask] = 0

    @torch.no_grad()
    def step(self, image: torch.Tensor, natural_language_embedding: torch.Tensor) -> Dict[str, torch.Tensor]:
        """One control step for all environments; returns the de-tokenised actions (views of static buffers: copy them if
        they must survive the next step)."""
        self.obs["image"].copy_(image, non_blocking=True)
        self.obs["natural_language_embedding"].copy_(natural_language_embedding, non_blocking=True)
        self.graph.replay()
        return self.actions
