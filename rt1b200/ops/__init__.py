"""Fused-op front end.

Every op has two implementations with the same signature:
  * ``rt1b200.ops.reference`` — plain PyTorch, the float32 oracle and the CPU path;
  * ``rt1b200.ops.kernels``   — hand-written sm_100a CUDA kernels (``rt1b200/csrc``) behind
    ``torch.autograd.Function`` wrappers, loaded through the C ABI in ``include/rt1b200.h``.

Dispatch rule: a CUDA **bf16** tensor takes the kernel path, anything else the reference path.  If a CUDA
bf16 tensor arrives and the native library cannot be loaded the call **raises** — there is no silent
fallback on a GPU box (set ``RT1B200_ALLOW_FALLBACK=1`` to override, e.g. to bisect a numerics issue).
``RT1B200_DISABLE_KERNELS=1`` forces the reference path everywhere.
"""
from __future__ import annotations

import os
from typing import Optional

import torch

from . import reference
from . import _lib

_DISABLED = os.environ.get("RT1B200_DISABLE_KERNELS", "0") == "1"
_ALLOW_FALLBACK = os.environ.get("RT1B200_ALLOW_FALLBACK", "0") == "1"
_force_reference = False


class force_reference:
    """Context manager: run the reference implementations regardless of device/dtype (used by tests)."""

    def __enter__(self):
        global _force_reference
        self._prev = _force_reference
        _force_reference = True

    def __exit__(self, *exc):
        global _force_reference
        _force_reference = self._prev


def kernels_available() -> bool:
    return _lib.available()


def use_kernels(t: torch.Tensor) -> bool:
    """True when ``t`` should go down the sm_100a kernel path."""
    if _DISABLED or _force_reference or not isinstance(t, torch.Tensor):
        return False
    if not (t.is_cuda and t.dtype == torch.bfloat16):
        return False
    if _lib.available():
        return True
    if _ALLOW_FALLBACK:
        return False
    raise RuntimeError(
        "rt1b200: CUDA bf16 tensor but the native kernel library is not loadable "
        f"({_lib.load_error()}); build it with `python -c 'import __graft_entry__ as g; g.build()'` "
        "or set RT1B200_ALLOW_FALLBACK=1 to run the PyTorch reference path.")


def _k():
    from . import kernels
    return kernels


# ----------------------------------------------------------------------------------------------- decoder
def linear(x, w, b=None, residual=None, act: Optional[str] = None, dropout_p: float = 0.0, wgrad=None, bgrad=None,
           notify=None):
    """``residual + dropout(act(x @ w.T + b))``; ``x [M,K]``, ``w [N,K]`` → ``[M,N]``.
    Kernel: tcgen05/TMEM GEMM fed by TMA with the bias/activation/dropout/residual epilogue fused.
    ``wgrad``/``bgrad``/``notify`` (kernel path only): explicit gradient destinations for fused operand views
    such as the concatenated q/k/v projection, and the parameters to announce as gradient-ready."""
    if use_kernels(x):
        return _k().linear(x, w, b, residual, act, dropout_p, wgrad=wgrad, bgrad=bgrad, notify=notify)
    seed = int(torch.randint(0, 2 ** 31 - 1, ()).item()) if dropout_p > 0 else 0
    return reference.linear(x, w, b, residual, act, dropout_p, seed)


def layer_norm(x, w, b, eps: float = 1e-5):
    if use_kernels(x):
        return _k().layer_norm(x, w, b, eps)
    return reference.layer_norm(x, w, b, eps)


def rt1_attention(qkv, mask, p_drop: float = 0.0):
    """Fused masked attention over the RT-1 token window. ``qkv [b,s,3,h,dk]`` → ``[b,s,h*dk]``."""
    if use_kernels(qkv):
        return _k().rt1_attention(qkv, mask, p_drop)
    seed = int(torch.randint(0, 2 ** 31 - 1, ()).item()) if p_drop > 0 else 0
    return reference.rt1_attention(qkv, mask, p_drop, seed)


def action_ce_loss(logits, labels, num_items: float):
    """256-bin action-token cross-entropy. ``logits [R,A,V]``, ``labels [R,A]`` → (loss ``[R]``, argmax ``[R,A]``)."""
    if use_kernels(logits):
        return _k().action_ce_loss(logits, labels, num_items)
    return reference.action_ce_loss(logits, labels, num_items)


def token_learner(feats, ln_w, ln_b, w1, b1, w2, b2, eps: float = 1e-5):
    if use_kernels(feats):
        return _k().token_learner(feats, ln_w, ln_b, w1, b1, w2, b2, eps)
    return reference.token_learner(feats, ln_w, ln_b, w1, b1, w2, b2, eps)


def preprocess_nhwc(image, shift, ratio: float = 0.07):
    if image.is_cuda and _lib.available() and not (_DISABLED or _force_reference):
        return _k().preprocess_nhwc(image, shift, ratio)
    return reference.preprocess_nhwc(image, shift, ratio)


from . import cnn  # noqa: E402  (NHWC image-tokenizer pipeline)

__all__ = ["reference", "cnn", "use_kernels", "kernels_available", "force_reference", "linear", "layer_norm",
           "rt1_attention", "action_ce_loss", "token_learner", "preprocess_nhwc"]
