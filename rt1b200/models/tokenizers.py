"""RT-1 image and action tokenizers.

Behavioural contract (reference ``pytorch_robotics_transformer/tokenizers/``):
  * ``RT1ImageTokenizer``   — ``image_tokenizer.py:31-84``: fold time into batch, FiLM-EfficientNet encoder
                               (no pooling), TokenLearner → ``[b, t, 8, 512]``; without TokenLearner the h*w
                               feature positions are the tokens.
  * ``RT1ActionTokenizer``  — ``action_tokenizer.py:68-159``: key order of the action space; Discrete → the value
                               itself; Box(n) → n tokens ``trunc((clamp(a)-low)/(high-low)*(vocab-1))``;
                               detokenise is the inverse affine map and sends Discrete tokens ``> n`` to 0.
  * ``batched_space_sampler`` / ``np_to_tensor`` — ``utils.py:8-25``.

Differences that are deliberate: no hard-coded ``'cuda'`` (bounds follow the input's device and are cached
per device), and no device→host synchronising assert on the token range.
"""
from __future__ import annotations

from typing import Dict, Optional, Union

import numpy as np
import torch
import torch.nn as nn

from .. import ops, spaces
from .film_efficientnet import EfficientNetEncoder
from .token_learner import TokenLearnerModule


class RT1ImageTokenizer(nn.Module):
    def __init__(self, embedding_output_dim: int = 512, use_token_learner: bool = False, num_tokens: int = 8,
                 weights: Optional[str] = None):
        super().__init__()
        self._tokenizer = EfficientNetEncoder(token_embedding_size=embedding_output_dim, weights=weights,
                                              early_film=True, pooling=False)
        self._use_token_learner = use_token_learner
        self._embedding_output_dim = embedding_output_dim
        if use_token_learner:
            self._num_tokens = num_tokens
            self._token_learner = TokenLearnerModule(inputs_channels=embedding_output_dim, num_tokens=num_tokens)

    @property
    def tokens_per_context_image(self) -> int:
        # 100 = 10x10 feature positions of a 300x300 input (reference ``image_tokenizer.py:45-50``)
        return self._num_tokens if self._use_token_learner else 100

    def forward(self, image: torch.Tensor, context: Optional[torch.Tensor] = None) -> torch.Tensor:
        """``image`` ``[b, t, 3, h, w]`` float, ``context`` ``[b, t, 512]`` → tokens ``[b, t, n_tok, emb]``."""
        b, t, c, h, w = image.shape
        image = image.reshape(b * t, c, h, w)
        if context is not None:
            context = context.reshape(b * t, -1)
        feats = self._tokenizer(image, context=context)                    # [bt, emb, h', w']
        if self._use_token_learner:
            tokens = self._token_learner(feats)
            return tokens.view(b, t, tokens.shape[1], -1)
        return feats.view(b, t, feats.shape[1], -1).transpose(2, 3)

    def forward_nhwc(self, image_nhwc: torch.Tensor, context: torch.Tensor, b: int, t: int) -> torch.Tensor:
        """sm_100a path: pre-processed NHWC bf16 frames ``[b*t, H, W, 8]`` → tokens ``[b, t, n_tok, emb]``."""
        feats = self._tokenizer.forward_nhwc(image_nhwc, context.reshape(b * t, -1))   # [bt,h',w',emb]
        if self._use_token_learner:
            tokens = self._token_learner.forward_nhwc(feats)
            return tokens.view(b, t, tokens.shape[1], -1)
        return feats.reshape(b, t, -1, feats.shape[-1])


class RT1ActionTokenizer:
    def __init__(self, action_space, vocab_size: int):
        self._action_space = action_space
        self._vocab_size = vocab_size
        self._action_order = list(action_space.keys())
        self._tokens_per_action = 0
        self._bounds_cache: Dict = {}
        for k in self._action_order:
            sp = action_space[k]
            if _is_discrete(sp):
                self._tokens_per_action += 1
            elif _is_box(sp):
                if len(sp.shape) != 1:
                    raise ValueError(f"Only action shapes with single dimension supported, got {sp.shape}")
                self._tokens_per_action += sp.shape[0]
            else:
                raise ValueError("action_space entries must be Discrete or Box")

    @property
    def tokens_per_action(self) -> int:
        return self._tokens_per_action

    @property
    def action_order(self):
        return list(self._action_order)

    def _bounds(self, key, device):
        ck = (key, str(device))
        if ck not in self._bounds_cache:
            sp = self._action_space[key]
            self._bounds_cache[ck] = (torch.as_tensor(np.asarray(sp.low), dtype=torch.float32, device=device),
                                      torch.as_tensor(np.asarray(sp.high), dtype=torch.float32, device=device))
        return self._bounds_cache[ck]

    # ---- sm_100a path: one kernel for the whole clamp / affine / truncate chain (csrc/action_tokens.cu) ----------------
    def _tables(self, device):
        """Per-token (low, high-or-n, is_discrete) tables on ``device`` (built once per device)."""
        ck = ("tables", str(device))
        if ck not in self._bounds_cache:
            low, high, disc = [], [], []
            for k in self._action_order:
                sp = self._action_space[k]
                if _is_discrete(sp):
                    low.append(0.0); high.append(float(sp.n)); disc.append(1)
                else:
                    lo = np.broadcast_to(np.asarray(sp.low, np.float32), sp.shape).reshape(-1)
                    hi = np.broadcast_to(np.asarray(sp.high, np.float32), sp.shape).reshape(-1)
                    low += lo.tolist(); high += hi.tolist(); disc += [0] * sp.shape[0]
            self._bounds_cache[ck] = (torch.tensor(low, dtype=torch.float32, device=device),
                                      torch.tensor(high, dtype=torch.float32, device=device),
                                      torch.tensor(disc, dtype=torch.uint8, device=device))
        return self._bounds_cache[ck]

    @staticmethod
    def _native(t: torch.Tensor) -> bool:
        from ..ops import _lib
        return t.is_cuda and _lib.available()

    def _tokenize_native(self, action: Dict[str, torch.Tensor]) -> torch.Tensor:
        from ..ops._lib import call, i32, i64, ptr, stream
        parts = []
        for k in self._action_order:
            a = action[k].to(torch.float32)
            parts.append(a.unsqueeze(-1) if _is_discrete(self._action_space[k]) else a)
        vals = torch.cat(parts, dim=-1).contiguous()
        low, high, disc = self._tables(vals.device)
        out = torch.empty(vals.shape, dtype=torch.int64, device=vals.device)
        A = vals.shape[-1]
        call("rt1_action_tokenize", ptr(vals), ptr(low), ptr(high), ptr(disc), ptr(out), i64(vals.numel() // A), i32(A),
             i32(self._vocab_size), stream())
        return out

    def _detokenize_native(self, action_tokens: torch.Tensor) -> Dict[str, torch.Tensor]:
        from ..ops._lib import call, i32, i64, ptr, stream
        tok = action_tokens.to(torch.int64).contiguous()
        low, high, disc = self._tables(tok.device)
        A = tok.shape[-1]
        vals = torch.empty(tok.shape, dtype=torch.float32, device=tok.device)
        call("rt1_action_detokenize", ptr(tok), ptr(low), ptr(high), ptr(disc), ptr(vals), i64(tok.numel() // A), i32(A),
             i32(self._vocab_size), stream())
        out: Dict[str, torch.Tensor] = {}
        i = 0
        for k in self._action_order:
            sp = self._action_space[k]
            if _is_discrete(sp):
                out[k] = vals[..., i].to(torch.int64)
                i += 1
            else:
                n = sp.shape[0]
                out[k] = vals[..., i:i + n]
                i += n
        return out

    def tokenize(self, action: Dict[str, torch.Tensor]) -> torch.Tensor:
        first = action[self._action_order[0]]
        if self._native(first) and first.numel() > 0:
            return self._tokenize_native(action)
        parts = []
        for k in self._action_order:
            a = action[k]
            sp = self._action_space[k]
            if _is_discrete(sp):
                parts.append(a.unsqueeze(-1).to(torch.int64))
            else:
                low, high = self._bounds(k, a.device)
                a = torch.clamp(a.to(torch.float32), low, high)
                tok = (a - low) / (high - low) * (self._vocab_size - 1)
                parts.append(tok.to(torch.int32).to(torch.int64))      # truncation, like the reference's int32 cast
        return torch.cat(parts, dim=-1)

    def detokenize(self, action_tokens: torch.Tensor) -> Dict[str, torch.Tensor]:
        if self._native(action_tokens) and action_tokens.numel() > 0 and action_tokens.shape[-1] == self._tokens_per_action:
            return self._detokenize_native(action_tokens)
        out: Dict[str, torch.Tensor] = {}
        i = 0
        for k in self._action_order:
            sp = self._action_space[k]
            if _is_discrete(sp):
                tok = action_tokens[..., i]
                out[k] = torch.where(tok > sp.n, torch.zeros_like(tok), tok)   # '>' (not '>='), as upstream
                i += 1
            else:
                n = sp.shape[0]
                low, high = self._bounds(k, action_tokens.device)
                tok = action_tokens[..., i:i + n].to(torch.float32)
                out[k] = tok / (self._vocab_size - 1) * (high - low) + low
                i += n
        return out


def _is_discrete(sp) -> bool:
    return type(sp).__name__ == "Discrete"


def _is_box(sp) -> bool:
    return type(sp).__name__ == "Box"


def batched_space_sampler(space, batch_size: int) -> Dict[str, np.ndarray]:
    """Sample ``batch_size`` elements of a Dict space and stack them per key (vector-env layout)."""
    samples = [space.sample() for _ in range(batch_size)]
    return {k: np.stack([np.asarray(s[k]) for s in samples], axis=0) for k in samples[0].keys()}


def np_to_tensor(sample_dict: Dict[str, Union[int, np.ndarray]], device=None) -> Dict[str, torch.Tensor]:
    return {k: torch.as_tensor(v).to(device) if device is not None else torch.as_tensor(v)
            for k, v in sample_dict.items()}
