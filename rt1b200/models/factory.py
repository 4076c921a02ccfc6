"""Model configurations named by BASELINE.json and a one-call constructor.

``RT1Config.language_table()`` is the reference training configuration (``distribute_train.py:28-55``):
8 layers, key_dim 128 x 8 heads, d_model 512, vocab 256, TokenLearner(8), 256x456 frames, T=6, action space
``terminate_episode: Discrete(2)``, ``action: Box(-0.1, 0.1, (2,))``.  ``RT1Config.scaled()`` is BASELINE
config 3 (24 layers, d_model 1024) used to stress the fused gradient all-reduce + AdamW.
"""
from __future__ import annotations

from collections import OrderedDict
from dataclasses import dataclass, replace
from typing import Optional

import numpy as np

from .. import spaces
from .transformer_network import TransformerNetwork


@dataclass
class RT1Config:
    height: int = 256
    width: int = 456
    seq_len: int = 6
    vocab_size: int = 256
    token_embedding_size: int = 512
    num_layers: int = 8
    layer_size: int = 128
    num_heads: int = 8
    feed_forward_size: int = 512
    dropout_rate: float = 0.1
    crop_size: int = 236
    use_token_learner: bool = True
    action_low: float = -0.1
    action_high: float = 0.1
    action_dim: int = 2
    weights: Optional[str] = None

    @classmethod
    def language_table(cls, **kw) -> "RT1Config":
        return replace(cls(), **kw)

    @classmethod
    def scaled(cls, **kw) -> "RT1Config":
        return replace(cls(num_layers=24, feed_forward_size=1024), **kw)

    @classmethod
    def tiny(cls, **kw) -> "RT1Config":
        """CPU-test sized: small frames, 2 layers."""
        return replace(cls(height=64, width=96, seq_len=3, num_layers=2, layer_size=32, num_heads=4), **kw)


def default_spaces(cfg: RT1Config):
    state_space = spaces.Dict({
        "image": spaces.Box(low=0.0, high=1.0, shape=(3, cfg.height, cfg.width), dtype=np.float32),
        "natural_language_embedding": spaces.Box(low=-np.inf, high=np.inf, shape=[512], dtype=np.float32),
    })
    action_space = spaces.Dict(OrderedDict([
        ("terminate_episode", spaces.Discrete(2)),
        ("action", spaces.Box(low=cfg.action_low, high=cfg.action_high, shape=(cfg.action_dim,), dtype=np.float32)),
    ]))
    return state_space, action_space


def build_rt1(cfg: Optional[RT1Config] = None, **kw) -> TransformerNetwork:
    cfg = cfg or RT1Config.language_table()
    state_space, action_space = default_spaces(cfg)
    return TransformerNetwork(
        input_tensor_space=state_space, output_tensor_space=action_space, vocab_size=cfg.vocab_size,
        token_embedding_size=cfg.token_embedding_size, num_layers=cfg.num_layers, layer_size=cfg.layer_size,
        num_heads=cfg.num_heads, feed_forward_size=cfg.feed_forward_size, dropout_rate=cfg.dropout_rate,
        time_sequence_length=cfg.seq_len, crop_size=cfg.crop_size, use_token_learner=cfg.use_token_learner,
        weights=cfg.weights, **kw)
