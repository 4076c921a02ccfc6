"""RT-1 causal decoder.

Behavioural contract (what the reference does, not how):
  * ``Transformer``            — reference ``pytorch_robotics_transformer/transformer.py:146-198``
  * ``_TransformerLayer``      — ``transformer.py:112-144`` (pre-LN, the "FFN" is ONE ``Linear(d, d)``,
                                  dropout only on that branch)
  * ``TF_MultiHeadAttention``  — ``transformer.py:29-79`` (Keras-style: q/k/v ``Linear(d, h*key_dim)`` with
                                  bias, out ``Linear(h*key_dim, d)``)
  * ``attention``              — ``transformer.py:82-109`` (mask==0 → -1e9, softmax, dropout on probs)

State-dict keys are identical to the reference (``_layers.N.{norm_1,attn.{q,k,v}_linear,attn.out,ff,norm_2}``,
``_token_emb``, ``_position_emb``, ``_output_tokens``).

Two execution paths share these parameter containers:
  * the *reference path* (plain PyTorch, any device/dtype) — the oracle for every kernel test;
  * the *sm_100a path* (``rt1b200.ops``: tcgen05 GEMMs with fused bias/residual epilogues, fused LayerNorm,
    one fused masked-attention kernel per layer) taken automatically for CUDA bf16 tensors.
"""
from __future__ import annotations

import math
from typing import List, Optional, Tuple

import torch
import torch.nn as nn
import torch.nn.functional as F

from .. import ops


def attention(q, k, v, key_dim, mask=None, dropout=None, return_attention_scores=False):
    """Reference-path scaled-dot-product attention with the RT-1 mask convention (0 = blocked)."""
    scores = torch.matmul(q, k.transpose(-2, -1)) / math.sqrt(key_dim)
    if mask is not None:
        m = mask.to(scores.device)
        while m.dim() < scores.dim():
            m = m.unsqueeze(0)
        scores = scores.masked_fill(m == 0, -1e9)
    probs = F.softmax(scores, dim=-1)
    if dropout is not None:
        probs = dropout(probs)
    out = torch.matmul(probs, v)
    return (out, probs) if return_attention_scores else out


class TF_MultiHeadAttention(nn.Module):
    def __init__(self, heads: int, d_model: int, key_dim: int, value_dim: Optional[int] = None,
                 dropout: float = 0.1, return_attention_scores: bool = False):
        super().__init__()
        self.d_model = d_model
        self.h = heads
        self.key_dim = key_dim
        self.value_dim = value_dim if value_dim else key_dim
        self.return_attention_scores = return_attention_scores
        self.q_linear = nn.Linear(d_model, heads * key_dim)
        self.k_linear = nn.Linear(d_model, heads * key_dim)
        self.v_linear = nn.Linear(d_model, heads * self.value_dim)
        self.dropout = nn.Dropout(dropout)
        self.out = nn.Linear(heads * self.value_dim, d_model)

    # Parameters the flat-buffer engine should lay out back to back so the three projections run as
    # one [3*h*dk, d] GEMM on the sm_100a path (see rt1b200.parallel.flat).
    def flat_groups(self):
        if self.value_dim != self.key_dim:
            return {}
        return {"_qkv_w": [self.q_linear.weight, self.k_linear.weight, self.v_linear.weight],
                "_qkv_b": [self.q_linear.bias, self.k_linear.bias, self.v_linear.bias]}

    def _fused_qkv(self):
        w = getattr(self, "_qkv_w", None)
        if w is not None and w.data_ptr() == self.q_linear.weight.data_ptr():
            return w, self._qkv_b
        return None, None

    def forward(self, q, k, v, mask=None, residual=None):
        """Returns ``out(attn(q,k,v))`` (+ ``residual`` if given — fused into the GEMM epilogue on sm_100a)."""
        bs, sl = q.shape[0], q.shape[1]
        if ops.use_kernels(q) and (q is k) and (k is v) and not self.return_attention_scores \
                and self.value_dim == self.key_dim:
            wqkv, bqkv = self._fused_qkv()
            x2 = q.reshape(bs * sl, self.d_model)
            if wqkv is not None:
                qkv = ops.linear(x2, wqkv, bqkv, wgrad=self._qkv_w_grad, bgrad=getattr(self, "_qkv_b_grad", None),
                                 notify=[self.q_linear.weight, self.k_linear.weight, self.v_linear.weight,
                                         self.q_linear.bias, self.k_linear.bias, self.v_linear.bias])   # [bs*sl, 3*h*dk]
            else:
                qkv = torch.cat([ops.linear(x2, self.q_linear.weight, self.q_linear.bias),
                                 ops.linear(x2, self.k_linear.weight, self.k_linear.bias),
                                 ops.linear(x2, self.v_linear.weight, self.v_linear.bias)], dim=1)
            p = self.dropout.p if self.training else 0.0
            ctx = ops.rt1_attention(qkv.view(bs, sl, 3, self.h, self.key_dim), mask, p)  # [bs, sl, h*dk]
            res2 = residual.reshape(bs * sl, self.d_model) if residual is not None else None
            out = ops.linear(ctx.reshape(bs * sl, self.h * self.key_dim), self.out.weight, self.out.bias,
                             residual=res2)
            return out.view(bs, sl, self.d_model)

        kk = self.k_linear(k).view(bs, -1, self.h, self.key_dim).transpose(1, 2)
        qq = self.q_linear(q).view(bs, -1, self.h, self.key_dim).transpose(1, 2)
        vv = self.v_linear(v).view(bs, -1, self.h, self.value_dim).transpose(1, 2)
        res = attention(qq, kk, vv, self.key_dim, mask, self.dropout, self.return_attention_scores)
        ctx, score = res if self.return_attention_scores else (res, None)
        concat = ctx.transpose(1, 2).contiguous().view(bs, -1, self.h * self.value_dim)
        out = self.out(concat)
        if residual is not None:
            out = residual + out
        return (out, score) if self.return_attention_scores else out


class _TransformerLayer(nn.Module):
    def __init__(self, layer_size: int = 4096, num_heads: int = 8, feed_forward_size: int = 512,
                 dropout_rate: float = 0.1, return_attention_scores: bool = False):
        super().__init__()
        self._return_attention_scores = return_attention_scores
        self.norm_1 = nn.LayerNorm(feed_forward_size)
        self.attn = TF_MultiHeadAttention(num_heads, feed_forward_size, layer_size, dropout=dropout_rate,
                                          return_attention_scores=return_attention_scores)
        self.ff = nn.Linear(feed_forward_size, feed_forward_size)
        self.norm_2 = nn.LayerNorm(feed_forward_size)
        self.dropout_1 = nn.Dropout(dropout_rate)

    def forward(self, x: torch.Tensor, mask: Optional[torch.Tensor]) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
        if ops.use_kernels(x) and not self._return_attention_scores:
            x1 = ops.layer_norm(x, self.norm_1.weight, self.norm_1.bias, self.norm_1.eps)
            x = self.attn(x1, x1, x1, mask=mask, residual=x)
            y = ops.layer_norm(x, self.norm_2.weight, self.norm_2.bias, self.norm_2.eps)
            p = self.dropout_1.p if self.training else 0.0
            shp = x.shape
            x = ops.linear(y.reshape(-1, shp[-1]), self.ff.weight, self.ff.bias,
                           residual=x.reshape(-1, shp[-1]), dropout_p=p).view(shp)
            return x, None
        x1 = self.norm_1(x)
        res = self.attn(x1, x1, x1, mask=mask)
        x1, score = res if self._return_attention_scores else (res, None)
        x = x + x1
        y = self.norm_2(x)
        x = x + self.dropout_1(self.ff(y))
        return x, score


class Transformer(nn.Module):
    def __init__(self, num_layers: int = 1, layer_size: int = 4096, num_heads: int = 8,
                 feed_forward_size: int = 512, dropout_rate: float = 0.1, vocab_size: int = 256,
                 input_token_emb_dim: int = 512, return_attention_scores: bool = False,
                 max_seq_len: int = 256):
        super().__init__()
        self._layers = nn.ModuleList([
            _TransformerLayer(layer_size=layer_size, num_heads=num_heads, feed_forward_size=feed_forward_size,
                              dropout_rate=dropout_rate, return_attention_scores=return_attention_scores)
            for _ in range(num_layers)])
        self._token_emb = nn.Linear(input_token_emb_dim, feed_forward_size)
        self._position_emb = nn.Embedding(max_seq_len, feed_forward_size)
        self._output_tokens = nn.Linear(feed_forward_size, vocab_size)

    def embed(self, inputs: torch.Tensor) -> torch.Tensor:
        bs, sl, _ = inputs.shape
        if ops.use_kernels(inputs):
            # positional table rows are the residual operand of the token-embedding GEMM epilogue
            pos = self._position_emb.weight[:sl].unsqueeze(0).expand(bs, sl, -1).reshape(bs * sl, -1)
            x = ops.linear(inputs.reshape(bs * sl, -1), self._token_emb.weight, self._token_emb.bias,
                           residual=pos)
            return x.view(bs, sl, -1)
        pos_ids = torch.arange(sl, dtype=torch.long, device=inputs.device)
        return self._token_emb(inputs) + self._position_emb(pos_ids).unsqueeze(0)

    def trunk(self, x: torch.Tensor, attention_mask) -> Tuple[torch.Tensor, List[torch.Tensor]]:
        from ..utils import trace
        scores = []
        for i, layer in enumerate(self._layers):
            with trace.range_(f"decoder/layer {i} fwd"):
                x, s = layer(x, mask=attention_mask)
            if s is not None:
                scores.append(s)
        return x, scores

    def head(self, x: torch.Tensor) -> torch.Tensor:
        if ops.use_kernels(x):
            shp = x.shape
            return ops.linear(x.reshape(-1, shp[-1]), self._output_tokens.weight,
                              self._output_tokens.bias).view(*shp[:-1], -1)
        return self._output_tokens(x)

    def forward(self, inputs: torch.Tensor, attention_mask):
        x = self.embed(inputs)
        x, scores = self.trunk(x, attention_mask)
        return self.head(x), scores
