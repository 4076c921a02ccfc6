"""CLIP text tower for the LAVA baseline's ``lang_encoder="clip"`` mode.

The reference builds it from an external package (``scenic.projects.baselines.clip.layers.TextEncoder`` with
``vocab_size=49408, features=512, num_heads=8, num_layers=12, out_features=512`` — ``language_table/train/networks/
lava.py:423-432``) and loads frozen pretrained weights; neither the package nor the checkpoint is vendored, and neither is
reachable offline.  This is the same architecture in PyTorch — CLIP ViT-B/32's text transformer: token + learned positional
embeddings, 12 pre-LayerNorm blocks with causal self-attention and a QuickGELU MLP, a final LayerNorm, the feature at the
end-of-text token (the highest token id of each sequence) and a bias-free projection — with a loader for the weight names of
the public CLIP checkpoints, so a user who has one can drop it in.  Randomly initialised it exercises the code path; it
carries no language knowledge.  The stored sentence embedding (``lang_encoder="clip_in_obs"``) remains the default.
"""
from __future__ import annotations

from typing import Dict

import torch
from torch import nn


class QuickGELU(nn.Module):
    def forward(self, x):
        return x * torch.sigmoid(1.702 * x)


class _Block(nn.Module):
    def __init__(self, width: int, heads: int):
        super().__init__()
        self.ln_1 = nn.LayerNorm(width)
        self.attn = nn.MultiheadAttention(width, heads, batch_first=True)
        self.ln_2 = nn.LayerNorm(width)
        self.mlp = nn.Sequential()
        self.mlp.add_module("c_fc", nn.Linear(width, 4 * width))
        self.mlp.add_module("gelu", QuickGELU())
        self.mlp.add_module("c_proj", nn.Linear(4 * width, width))

    def forward(self, x, mask):
        h = self.ln_1(x)
        x = x + self.attn(h, h, h, need_weights=False, attn_mask=mask)[0]
        return x + self.mlp(self.ln_2(x))


class ClipTextEncoder(nn.Module):
    """``tokens int64 [b, context]`` (CLIP BPE ids, zero padded, one end-of-text id = the largest id) → ``[b, out_features]``."""

    def __init__(self, vocab_size: int = 49408, features: int = 512, num_heads: int = 8, num_layers: int = 12,
                 out_features: int = 512, context_length: int = 77):
        super().__init__()
        self.context_length = context_length
        self.token_embedding = nn.Embedding(vocab_size, features)
        self.positional_embedding = nn.Parameter(torch.empty(context_length, features))
        self.resblocks = nn.ModuleList([_Block(features, num_heads) for _ in range(num_layers)])
        self.ln_final = nn.LayerNorm(features)
        self.text_projection = nn.Parameter(torch.empty(features, out_features))
        self.register_buffer("_mask", torch.full((context_length, context_length), float("-inf")).triu_(1), persistent=False)
        self._init(features, num_layers)

    def _init(self, width: int, layers: int):
        # the initialisation of the public CLIP code: small embeddings, depth-scaled projections
        nn.init.normal_(self.token_embedding.weight, std=0.02)
        nn.init.normal_(self.positional_embedding, std=0.01)
        proj_std, attn_std, fc_std = (width ** -0.5) * ((2 * layers) ** -0.5), width ** -0.5, (2 * width) ** -0.5
        for blk in self.resblocks:
            nn.init.normal_(blk.attn.in_proj_weight, std=attn_std)
            nn.init.normal_(blk.attn.out_proj.weight, std=proj_std)
            nn.init.normal_(blk.mlp.c_fc.weight, std=fc_std)
            nn.init.normal_(blk.mlp.c_proj.weight, std=proj_std)
        nn.init.normal_(self.text_projection, std=width ** -0.5)

    def forward(self, tokens: torch.Tensor) -> torch.Tensor:
        b, n = tokens.shape
        if n > self.context_length:
            raise ValueError(f"{n} tokens exceed the context length {self.context_length}")
        x = self.token_embedding(tokens) + self.positional_embedding[:n]
        mask = self._mask[:n, :n]
        for blk in self.resblocks:
            x = blk(x, mask)
        x = self.ln_final(x)
        eot = tokens.argmax(dim=-1)                      # the end-of-text id is the largest id of the vocabulary
        return x[torch.arange(b, device=x.device), eot] @ self.text_projection

    def freeze(self) -> "ClipTextEncoder":
        """The reference keeps the tower frozen (its parameters are not in the optimiser's tree)."""
        for p in self.parameters():
            p.requires_grad_(False)
        return self.eval()


def load_clip_text_weights(module: ClipTextEncoder, state: Dict[str, torch.Tensor]) -> ClipTextEncoder:
    """Copy the text-tower entries of a public CLIP ``state_dict`` (keys ``token_embedding.weight``,
    ``positional_embedding``, ``transformer.resblocks.<i>.…``, ``ln_final.*``, ``text_projection``; a ``visual.*`` half and
    ``logit_scale`` are ignored) into ``module``.  Raises if a text-tower entry is missing or has another shape."""
    own = module.state_dict()
    renamed = {}
    for k, v in state.items():
        if k.startswith("visual.") or k in ("logit_scale", "input_resolution", "context_length", "vocab_size"):
            continue
        k2 = k[len("transformer."):] if k.startswith("transformer.resblocks.") else k
        renamed[k2] = v
    missing = [k for k in own if k not in renamed]
    if missing:
        raise KeyError(f"CLIP checkpoint lacks text-tower entries: {missing[:5]}{' ...' if len(missing) > 5 else ''}")
    for k, dst in own.items():
        src = renamed[k]
        if tuple(src.shape) != tuple(dst.shape):
            raise ValueError(f"{k}: checkpoint shape {tuple(src.shape)} != {tuple(dst.shape)}")
    module.load_state_dict({k: renamed[k].to(own[k].dtype) for k in own})
    return module
