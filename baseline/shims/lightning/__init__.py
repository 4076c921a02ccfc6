"""Import shim for ``lightning`` (absent offline).  Only what ``distribute_train.py`` touches at import time and
inside ``training_step``: ``LightningModule`` is an ``nn.Module`` whose ``log`` is a no-op; ``Trainer`` is not
provided as a training loop — the benchmark harness drives ``training_step`` under torch DDP + Adam, which is
what Lightning's ``DDPStrategy(find_unused_parameters=True)`` does."""
import torch.nn as nn


class LightningModule(nn.Module):
    def log(self, *args, **kwargs):
        return None

    def log_dict(self, *args, **kwargs):
        return None

    def save_hyperparameters(self, *args, **kwargs):
        return None


class Trainer:  # pragma: no cover
    def __init__(self, *a, **k):
        raise RuntimeError("lightning shim: Trainer is not available; see baseline/run_reference.py")
