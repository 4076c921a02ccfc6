"""TokenLearner attention pooling (Ryoo et al. 2021) as used by RT-1.

Behavioural contract: reference ``pytorch_robotics_transformer/tokenizers/token_learner.py:26-95`` —
LayerNorm over channels → 1x1 conv C→64 → GELU(tanh) → 1x1 conv 64→``num_tokens`` → softmax over the h*w
positions → weighted sum of the *un-normalised* input features.  Keys: ``layerNorm``, ``conv1``, ``conv2``.

On the sm_100a path the whole module is ONE kernel per direction (``rt1b200.ops.token_learner``): the
120x512 feature map of a frame fits in shared memory, so LN, both projections, the softmax and the pooling
read HBM exactly once.
"""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

from .. import ops


class TokenLearnerModule(nn.Module):
    def __init__(self, inputs_channels: int, num_tokens: int, bottleneck_dim: int = 64, dropout_rate: float = 0.0):
        super().__init__()
        self.layerNorm = nn.LayerNorm(inputs_channels)
        self.conv1 = nn.Conv2d(inputs_channels, bottleneck_dim, 1)
        self.gelu1 = nn.GELU(approximate="tanh")
        self.dropout1 = nn.Dropout2d(dropout_rate) if dropout_rate > 0 else nn.Identity()
        self.conv2 = nn.Conv2d(bottleneck_dim, num_tokens, 1)
        self.dropout2 = nn.Dropout2d(dropout_rate) if dropout_rate > 0 else nn.Identity()
        self._dropout_rate = dropout_rate

    def forward(self, inputs: torch.Tensor) -> torch.Tensor:
        """``inputs`` NCHW ``[n, c, h, w]`` → ``[n, num_tokens, c]``."""
        n, c, h, w = inputs.shape
        feats = inputs.permute(0, 2, 3, 1)                                # [n,h,w,c]
        x = self.layerNorm(feats).permute(0, 3, 1, 2)
        x = self.dropout1(self.gelu1(self.conv1(x)))
        x = self.dropout2(self.conv2(x))                                  # [n,T,h,w]
        weights = F.softmax(x.reshape(n, x.shape[1], h * w), dim=-1)
        return torch.bmm(weights, feats.reshape(n, h * w, c))

    def forward_nhwc(self, feats: torch.Tensor) -> torch.Tensor:
        """``feats`` NHWC ``[n, h, w, c]`` (sm_100a path, bf16) → ``[n, num_tokens, c]``."""
        n, h, w, c = feats.shape
        if ops.use_kernels(feats) and self._dropout_rate == 0:
            return ops.token_learner(feats.reshape(n, h * w, c), self.layerNorm.weight, self.layerNorm.bias,
                                     self.conv1.weight, self.conv1.bias, self.conv2.weight, self.conv2.bias,
                                     self.layerNorm.eps)
        return self.forward(feats.permute(0, 3, 1, 2))
