"""Plain-PyTorch statements of every fused op in :mod:`rt1b200.ops`.

These are the numerics oracles: each sm_100a kernel is tested against the function of the same name here,
evaluated in float32.  They also serve as the execution path on CPU / non-bf16 tensors.  Shapes and
semantics are documented once, on the dispatching wrappers in ``rt1b200/ops/__init__.py``.
"""
from __future__ import annotations

import math
from typing import Optional, Tuple

import torch
import torch.nn.functional as F

# ----------------------------------------------------------------------------- counter-based dropout RNG
# The kernels draw dropout bits from a stateless integer hash of (seed, element index) so the backward pass
# can regenerate the mask instead of storing it.  Same hash here, in int64 arithmetic masked to 32 bits.
_M32 = 0xFFFFFFFF


def hash_u32(idx: torch.Tensor, seed: int) -> torch.Tensor:
    """lowbias32-style avalanche hash of ``idx + seed * 0x9E3779B9`` (all arithmetic mod 2^32)."""
    x = (idx.to(torch.int64) + (int(seed) * 0x9E3779B9 & _M32)) & _M32
    x = (x ^ (x >> 16)) & _M32
    x = (x * 0x7FEB352D) & _M32
    x = (x ^ (x >> 15)) & _M32
    x = (x * 0x846CA68B) & _M32
    x = (x ^ (x >> 16)) & _M32
    return x


def dropout_keep_mask(numel: int, p: float, seed: int, device=None) -> torch.Tensor:
    """Boolean keep-mask for a flat tensor of ``numel`` elements: keep iff hash >= p * 2^32."""
    idx = torch.arange(numel, device=device, dtype=torch.int64)
    thresh = min(int(p * 4294967296.0), _M32)
    return hash_u32(idx, seed) >= thresh


def dropout(x: torch.Tensor, p: float, seed: int) -> torch.Tensor:
    if p <= 0.0:
        return x
    keep = dropout_keep_mask(x.numel(), p, seed, x.device).view(x.shape)
    return x * keep.to(x.dtype) / (1.0 - p)


# ----------------------------------------------------------------------------- decoder ops
def linear(x, w, b=None, residual=None, act: Optional[str] = None, dropout_p: float = 0.0, seed: int = 0):
    y = F.linear(x, w, b)
    if act == "silu":
        y = F.silu(y)
    elif act == "gelu_tanh":
        y = F.gelu(y, approximate="tanh")
    elif act is not None:
        raise ValueError(act)
    if dropout_p > 0.0:
        y = dropout(y, dropout_p, seed)
    if residual is not None:
        y = residual + y
    return y


def layer_norm(x, w, b, eps: float = 1e-5):
    return F.layer_norm(x, (x.shape[-1],), w, b, eps)


def rt1_attention(qkv, mask, p_drop: float = 0.0, seed: int = 0):
    """``qkv`` ``[b, s, 3, h, dk]`` → ``[b, s, h*dk]``; ``mask`` ``[s, s]`` with 0 = blocked (→ -1e9)."""
    b, s, _, h, dk = qkv.shape
    q, k, v = (qkv[:, :, i].transpose(1, 2) for i in range(3))           # [b,h,s,dk]
    scores = torch.matmul(q, k.transpose(-2, -1)) / math.sqrt(dk)
    if mask is not None:
        scores = scores.masked_fill(mask.to(scores.device)[None, None] == 0, -1e9)
    probs = F.softmax(scores.float(), dim=-1).to(qkv.dtype)
    if p_drop > 0.0:
        probs = dropout(probs, p_drop, seed)
    out = torch.matmul(probs, v)                                          # [b,h,s,dk]
    return out.transpose(1, 2).reshape(b, s, h * dk)


def action_ce_loss(logits, labels, num_items: float) -> Tuple[torch.Tensor, torch.Tensor]:
    """``logits`` ``[R, A, V]``, ``labels`` ``[R, A]`` → (``mean_A(CE)/num_items`` ``[R]``, argmax ``[R, A]``)."""
    R, A, V = logits.shape
    ce = F.cross_entropy(logits.reshape(R * A, V).float(), labels.reshape(R * A), reduction="none").view(R, A)
    return ce.mean(dim=-1) / num_items, logits.argmax(dim=-1)


def token_learner(feats, ln_w, ln_b, w1, b1, w2, b2, eps: float = 1e-5):
    """``feats`` ``[n, P, C]`` → ``[n, T, C]``; ``w1`` ``[64, C(,1,1)]``, ``w2`` ``[T, 64(,1,1)]``."""
    n, P, C = feats.shape
    x = F.layer_norm(feats, (C,), ln_w, ln_b, eps)
    hdn = F.gelu(F.linear(x, w1.reshape(w1.shape[0], -1), b1), approximate="tanh")
    logit = F.linear(hdn, w2.reshape(w2.shape[0], -1), b2)                # [n,P,T]
    wmap = F.softmax(logit.float(), dim=1).to(feats.dtype)                # over positions
    return torch.bmm(wmap.transpose(1, 2), feats)


def preprocess_nhwc(image, shift, ratio: float = 0.07, pad_to: int = 8, dtype=torch.bfloat16):
    """``image`` ``[N,3,H,W]`` (uint8 0..255 or float 0..1) → ``[N,H,W,pad_to]`` with the pad-shift crop."""
    from ..models.preprocessors import convert_dtype_and_crop_images
    x = convert_dtype_and_crop_images(image, ratio=ratio, shift=shift)    # [N,3,H,W] f32
    x = x.permute(0, 2, 3, 1)
    if pad_to > x.shape[-1]:
        x = F.pad(x, (0, pad_to - x.shape[-1]))
    return x.contiguous().to(dtype)


# ----------------------------------------------------------------------------- optimizer
def adamw_step(p, g, m, v, lr, beta1, beta2, eps, weight_decay, step, grad_scale: float = 1.0):
    """One decoupled-weight-decay Adam step on fp32 tensors, in place; identical to ``torch.optim.Adam`` when
    ``weight_decay == 0`` (the reference recipe, ``distribute_train.py:100``)."""
    g = g.float() * grad_scale
    if weight_decay != 0.0:
        p.mul_(1.0 - lr * weight_decay)
    m.mul_(beta1).add_(g, alpha=1.0 - beta1)
    v.mul_(beta2).addcmul_(g, g, value=1.0 - beta2)
    bc1 = 1.0 - beta1 ** step
    bc2 = 1.0 - beta2 ** step
    denom = (v.sqrt() / math.sqrt(bc2)).add_(eps)
    p.addcdiv_(m, denom, value=-lr / bc1)
    return p
