"""Behaviour-cloning agent for the LAVA baseline (reference ``language_table/train/bc.py:51-234``): MSE between the
predicted and the (normalised) dataset action, Adam with eps 1e-7, optional frozen parameter groups, optional gradient
averaging over a ``torch.distributed`` group (the reference's ``jax.lax.pmean`` over the ``"batch"`` pmap axis)."""
import dataclasses
import math
from typing import Any, Callable, Dict, Optional, Sequence, Tuple

import torch
import torch.nn as nn


@dataclasses.dataclass
class TrainState:
    """What a checkpoint holds (reference ``bc.py:32-38``).  The agent owns the live module and optimiser; a
    ``TrainState`` is a view of them — ``params`` the trainable tensors, ``batch_stats`` the BatchNorm buffers,
    ``opt_state`` the optimiser's state dict."""
    step: int
    params: Any
    opt_state: Any
    batch_stats: Any
    norm_info: Any

    def replace(self, **changes) -> "TrainState":
        return dataclasses.replace(self, **changes)


@dataclasses.dataclass
class TrainMetrics:
    """Loss mean / standard deviation and the last learning rate over the steps merged so far (reference
    ``bc.py:41-47``: ``Average``, ``Std`` and ``LastValue`` of a clu metrics collection)."""
    count: int = 0
    loss_sum: float = 0.0
    loss_sumsq: float = 0.0
    last_learning_rate: float = 0.0

    @classmethod
    def empty(cls) -> "TrainMetrics":
        return cls()

    @classmethod
    def single_from_model_output(cls, loss, learning_rate=0.0, **unused) -> "TrainMetrics":
        loss = torch.as_tensor(loss, dtype=torch.float64).reshape(-1)
        lr = torch.as_tensor(learning_rate, dtype=torch.float64).reshape(-1)[-1]
        return cls(int(loss.numel()), float(loss.sum()), float((loss * loss).sum()), float(lr))

    gather_from_model_output = single_from_model_output

    def merge(self, other: "TrainMetrics") -> "TrainMetrics":
        return TrainMetrics(self.count + other.count, self.loss_sum + other.loss_sum, self.loss_sumsq + other.loss_sumsq,
                            other.last_learning_rate if other.count else self.last_learning_rate)

    def compute(self) -> Dict[str, float]:
        n = max(self.count, 1)
        mean = self.loss_sum / n
        return {"learning_rate": self.last_learning_rate, "loss": mean,
                "loss_std": math.sqrt(max(self.loss_sumsq / n - mean * mean, 0.0))}


def flattened_traversal(fn: Callable[[Tuple[str, ...], Any], Any]):
    """``fn(path, leaf)`` applied over a nested dict, the result in the same nesting (reference ``bc.py:237-245``; used
    there to label parameters "adam" or "zero" for the frozen groups)."""
    def mask(tree, _prefix=()):
        return {k: mask(v, _prefix + (k,)) if isinstance(v, dict) else fn(_prefix + (k,), v) for k, v in tree.items()}
    return mask


def _nest(flat: Dict[str, Any], sep: str = ".") -> Dict:
    out: Dict = {}
    for key, v in flat.items():
        node = out
        parts = key.split(sep)
        for part in parts[:-1]:
            node = node.setdefault(part, {})
        node[parts[-1]] = v
    return out


class BCAgent:
    def __init__(self, model: nn.Module, sequence_length: int, learning_rate: float, observation_statistics: Dict,
                 action_statistics: Dict, action_min, action_max, pretrained_checkpoints: Optional[Sequence] = None,
                 freeze_keys: Optional[Sequence[str]] = None, process_group=None):
        self.model, self.sequence_length, self.learning_rate = model, sequence_length, learning_rate
        self.pretrained_checkpoints = list(pretrained_checkpoints or [])
        self.loaded_pretrained = self._load_pretrained()
        self.norm_info = {"observation_statistics": observation_statistics, "action_statistics": action_statistics,
                          "action_min": action_min, "action_max": action_max}
        self.freeze_keys = list(freeze_keys or [])
        for name, p in model.named_parameters():
            if any(k in name for k in self.freeze_keys):
                p.requires_grad_(False)
        self.optimizer = torch.optim.Adam([p for p in model.parameters() if p.requires_grad], lr=learning_rate, eps=1e-7)
        self.step = 0
        self.group = process_group

    def _load_pretrained(self):
        """``pretrained_checkpoints``: ``[(path, [(prefix in the file, prefix in this model), ...]), ...]``.  Every entry of
        the file whose name starts with a source prefix is copied into the model entry of the renamed key when that
        exists with the same shape (reference ``bc.py:92-108``; a file is a ``torch.save``-d state dict, or a training
        checkpoint of this package, whose ``params`` are used).  Returns the ``(from, to)`` pairs that were copied."""
        loaded = []
        own = self.model.state_dict()
        for path, replacements in self.pretrained_checkpoints:
            sd = torch.load(path, map_location="cpu", weights_only=False)
            sd = sd.get("params", sd) if isinstance(sd, dict) else sd
            for src, dst in replacements:
                for key, value in sd.items():
                    if not key.startswith(src):
                        continue
                    target = key.replace(src, dst, 1)
                    if target in own and tuple(own[target].shape) == tuple(value.shape):
                        own[target].copy_(value.to(own[target].dtype))
                        loaded.append((key, target))
        return loaded

    def _split_state(self):
        params = {k: v for k, v in self.model.named_parameters()}
        stats = {k: v for k, v in self.model.named_buffers()}
        return params, stats

    def create_train_state(self, batch=None, rng=None):
        """``(TrainState, initial TrainMetrics)`` of the agent as it stands (reference ``bc.py:84-176``).  The module is
        built — and pretrained entries copied, frozen groups excluded from the optimiser — in ``__init__``; ``batch`` and
        ``rng`` are accepted for the reference's call and unused: a PyTorch module needs no sample input to initialise."""
        params, stats = self._split_state()
        return (TrainState(step=self.step, params=params, opt_state=self.optimizer.state_dict(), batch_stats=stats,
                           norm_info=self.norm_info), TrainMetrics.empty())

    def frozen_labels(self) -> Dict:
        """``"zero"`` / ``"adam"`` per parameter in the nesting of the module tree — the reference's optimiser labelling."""
        return flattened_traversal(lambda path, p: "adam" if p.requires_grad else "zero")(_nest(dict(self.model.named_parameters())))

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
