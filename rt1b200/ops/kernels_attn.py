"""Autograd wrappers for the fused masked-attention and TokenLearner kernels (``csrc/attention.cu``,
``csrc/token_learner.cu``)."""
from __future__ import annotations

import ctypes
from typing import Optional

import torch
from torch.autograd import Function

from . import _lib
from ._lib import call, f32, i32, i64, ptr, stream

BF16 = torch.bfloat16


def _u32(v):
    return ctypes.c_uint(int(v) & 0xFFFFFFFF)


class _Attention(Function):
    @staticmethod
    def forward(ctx, qkv, mask, p_drop, seed):
        b, s, three, h, dk = qkv.shape
        assert three == 3
        qkv = qkv if qkv.is_contiguous() else qkv.contiguous()
        if mask is not None:
            mask = mask.to(device=qkv.device, dtype=torch.uint8).contiguous()
        out = torch.empty(b, s, h * dk, dtype=BF16, device=qkv.device)
        call("rt1_attention_fwd", ptr(qkv), ptr(mask), ptr(out), i32(b), i32(s), i32(h), i32(dk), f32(p_drop), _u32(seed),
             stream())
        ctx.save_for_backward(qkv, mask)
        ctx.meta = (p_drop, seed)
        return out

    @staticmethod
    def backward(ctx, dout):
        qkv, mask = ctx.saved_tensors
        p_drop, seed = ctx.meta
        b, s, _, h, dk = qkv.shape
        dout = dout if dout.is_contiguous() else dout.contiguous()
        dqkv = torch.empty_like(qkv)
        call("rt1_attention_bwd", ptr(qkv), ptr(mask), ptr(dout), ptr(dqkv), i32(b), i32(s), i32(h), i32(dk), f32(p_drop),
             _u32(seed), stream())
        return dqkv, None, None, None


def rt1_attention(qkv, mask, p_drop: float = 0.0, seed: Optional[int] = None):
    from . import kernels
    if seed is None:
        seed = kernels.next_seed() if p_drop > 0 else 0
    return _Attention.apply(qkv, mask, float(p_drop), int(seed))


class _TokenLearner(Function):
    """TokenLearner as LayerNorm kernel → tcgen05 GEMM (C → 64, all frames at once) → per-frame pooling kernel; the
    backward mirrors it (pooling backward → two GEMMs for dLN(x) and dW1 → LayerNorm backward → direct pooling term).
    Envelope of the pooling kernels: C a multiple of 64, at most 8 learned tokens, at most 512 positions per frame."""

    @staticmethod
    def forward(ctx, feats, ln_w, ln_b, w1, b1, w2, b2, eps, params):
        from .kernels import gemm
        n, P, C = feats.shape
        T = w2.shape[0]
        feats = feats if feats.is_contiguous() else feats.contiguous()
        w1m = w1.reshape(w1.shape[0], -1)
        w2m = w2.reshape(T, -1)
        assert w1m.shape == (64, C) and w2m.shape == (T, 64), "TokenLearner kernel expects a 64-wide bottleneck"
        tokens = torch.empty(n, T, C, dtype=BF16, device=feats.device)
        if C % 64 != 0 or P > 512 or T > 8:
            raise ValueError(f"TokenLearner kernels support C % 64 == 0, <= 8 tokens and <= 512 positions per frame; got C={C}, "
                             f"T={T}, P={P}")
        ctx.meta = (eps, params)
        rows = n * P
        x2 = feats.view(rows, C)
        xn = torch.empty_like(x2)
        mean = torch.empty(rows, dtype=torch.float32, device=feats.device)
        rstd = torch.empty_like(mean)
        call("rt1_layernorm_fwd", ptr(x2), ptr(ln_w), ptr(ln_b), ptr(xn), ptr(mean), ptr(rstd), i32(rows), i32(C), f32(eps), stream())
        hpre = torch.empty(rows, 64, dtype=BF16, device=feats.device)
        gemm(xn, w1m, hpre, rows, 64, C, bias=b1)
        wmap = torch.empty(n, P, T, dtype=torch.float32, device=feats.device)
        call("rt1_tl_pool_fwd", ptr(feats), ptr(hpre), ptr(w2m), ptr(b2), ptr(tokens), i64(T * C), ptr(wmap), i32(n), i32(P), i32(C),
             i32(T), stream())
        ctx.save_for_backward(feats, ln_w, w1m, w2m, xn, mean, rstd, hpre, wmap)
        return tokens

    @staticmethod
    def backward(ctx, dtok):
        from .kernels import _grad_slot, _ready, gemm
        eps, params = ctx.meta
        dtok = dtok if dtok.is_contiguous() else dtok.contiguous()
        g = [_grad_slot(p) for p in params]     # order: ln_w, ln_b, w1, b1, w2, b2
        from .kernels_cnn import _wgrad_gemm
        feats, ln_w, w1m, w2m, xn, mean, rstd, hpre, wmap = ctx.saved_tensors
        n, P, C = feats.shape
        T = w2m.shape[0]
        rows = n * P
        dev = feats.device
        dhpre = torch.empty(rows, 64, dtype=BF16, device=dev)
        scratch = torch.empty(T * 64 + T, dtype=torch.float32, device=dev)
        call("rt1_tl_pool_bwd", ptr(feats), ptr(hpre), ptr(wmap), ptr(w2m), ptr(dtok), i64(T * C), ptr(dhpre), ptr(scratch), i32(n),
             i32(P), i32(C), i32(T), stream())
        call("rt1_cast_f32_bf16", ptr(scratch), ptr(g[4]), i64(T * 64), stream())                      # dW2
        call("rt1_cast_f32_bf16", ptr(scratch[T * 64:]), ptr(g[5]), i64(T), stream())                  # db2
        call("rt1_colsum_bf16", ptr(dhpre), i64(64), i32(rows), i32(64), ptr(g[3]), stream())          # db1
        _wgrad_gemm(dhpre, xn, params[2], 64, C, rows)                                                 # dW1 = dhpre^T · LN(x)
        dxn = torch.empty(rows, C, dtype=BF16, device=dev)
        gemm(dhpre, w1m, dxn, rows, C, 64, b_mn=True)                                                  # dLN(x) = dhpre · W1
        dfeats = torch.empty_like(feats)
        n_scratch = min(148, (rows + 7) // 8) * 2 * C
        ln_scratch = torch.empty(n_scratch, dtype=torch.float32, device=dev)
        call("rt1_layernorm_bwd", ptr(feats.view(rows, C)), ptr(ln_w), ptr(dxn), ptr(mean), ptr(rstd), ptr(dfeats), ptr(g[0]), ptr(g[1]),
             ptr(ln_scratch), i32(rows), i32(C), stream())
        call("rt1_tl_direct_bwd", ptr(wmap), ptr(dtok), i64(T * C), ptr(dfeats), i32(n), i32(P), i32(C), i32(T), stream())
        _ready(params)
        return (dfeats,) + (None,) * 8


def token_learner(feats, ln_w, ln_b, w1, b1, w2, b2, eps: float = 1e-5):
    params = (ln_w, ln_b, w1, b1, w2, b2)
    return _TokenLearner.apply(feats, ln_w.detach(), ln_b.detach(), w1.detach(), b1.detach(), w2.detach(), b2.detach(),
                               float(eps), params)
