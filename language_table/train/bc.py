"""Behaviour-cloning agent for the LAVA baseline (reference ``language_table/train/bc.py:51-234``): MSE between the
predicted and the (normalised) dataset action, Adam with eps 1e-7, optional frozen parameter groups, optional gradient
averaging over a ``torch.distributed`` group (the reference's ``jax.lax.pmean`` over the ``"batch"`` pmap axis)."""
from typing import Dict, Optional, Sequence

import torch
import torch.nn as nn


class BCAgent:
    def __init__(self, model: nn.Module, sequence_length: int, learning_rate: float, observation_statistics: Dict,
                 action_statistics: Dict, action_min, action_max, freeze_keys: Optional[Sequence[str]] = None,
                 process_group=None):
        self.model, self.sequence_length, self.learning_rate = model, sequence_length, learning_rate
        self.norm_info = {"observation_statistics": observation_statistics, "action_statistics": action_statistics,
                          "action_min": action_min, "action_max": action_max}
        self.freeze_keys = list(freeze_keys or [])
        for name, p in model.named_parameters():
            if any(k in name for k in self.freeze_keys):
                p.requires_grad_(False)
        self.optimizer = torch.optim.Adam([p for p in model.parameters() if p.requires_grad], lr=learning_rate, eps=1e-7)
        self.step = 0
        self.group = process_group

    def bc_loss(self, batch) -> torch.Tensor:
        pred = self.model(batch["observation"])
        return torch.mean((pred - batch["action"]) ** 2)

    def train(self, batch) -> Dict[str, float]:
        self.model.train()
        self.optimizer.zero_grad(set_to_none=True)
        loss = self.bc_loss(batch)
        loss.backward()
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized() and dist.get_world_size(self.group) > 1:
            world = dist.get_world_size(self.group)
            for p in self.model.parameters():
                if p.grad is not None:
                    dist.all_reduce(p.grad, group=self.group)
                    p.grad.div_(world)
            loss = loss.detach().clone()
            dist.all_reduce(loss, group=self.group)
            loss = loss / world
        self.optimizer.step()
        self.step += 1
        return {"loss": float(loss.detach()), "learning_rate": self.learning_rate}

    @torch.no_grad()
    def act(self, observation) -> torch.Tensor:
        self.model.eval()
        return self.model(observation)

    def state_dict(self) -> Dict:
        return {"step": self.step, "params": self.model.state_dict(), "opt_state": self.optimizer.state_dict(),
                "norm_info": self.norm_info}

    def load_state_dict(self, sd: Dict):
        self.step = int(sd["step"])
        self.model.load_state_dict(sd["params"])
        self.optimizer.load_state_dict(sd["opt_state"])
        self.norm_info = sd["norm_info"]
