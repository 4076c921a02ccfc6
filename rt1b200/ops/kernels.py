"""``torch.autograd.Function`` wrappers over the sm_100a kernels in ``librt1b200.so``.

Conventions
-----------
* activations and parameters are bf16, contiguous; reductions/statistics are fp32;
* **parameter gradients are written by the backward kernels straight into ``param.grad``** (which the
  data-parallel engine has pre-pointed into its flat, NVLink-visible gradient buffer) and ``None`` is returned
  to autograd for them — no per-parameter accumulate kernels, no extra copy.  ``GRAD_READY`` (set by the engine)
  is called with the parameters whose gradient has just been produced, which is what drives the
  bucket-by-bucket overlap of the gradient exchange with the rest of backward;
* every call goes through the C ABI declared in ``include/rt1b200.h`` on the current CUDA stream.
"""
from __future__ import annotations

import math
from typing import Callable, List, Optional, Sequence

import torch
from torch.autograd import Function

from . import _lib
from ._lib import call, f32, i32, i64, ptr, stream

BF16 = torch.bfloat16

# set by DataParallelEngine: called with a list of nn.Parameter whose .grad was just written by a kernel
GRAD_READY: Optional[Callable[[Sequence[torch.nn.Parameter]], None]] = None
_seed_counter = [0x1234567]


def next_seed() -> int:
    """Deterministic per-call dropout seed stream (reproducible under torch.manual_seed via set_seed)."""
    _seed_counter[0] = (_seed_counter[0] * 1103515245 + 12345) & 0x7FFFFFFF
    return _seed_counter[0]


def set_seed(seed: int):
    _seed_counter[0] = seed & 0x7FFFFFFF


def _grad_slot(p: torch.Tensor) -> torch.Tensor:
    if p.grad is None:
        p.grad = torch.zeros_like(p)
    return p.grad


def _ready(params):
    if GRAD_READY is not None and params:
        GRAD_READY([p for p in params if p is not None])


class WgradLane:
    """Weight-gradient kernels on a SECOND stream.

    In a backward pass only the data gradients form the dependency chain; the weight gradients (split-K GEMMs of the 1x1
    convolutions and linears, depthwise wgrad, bias column sums) are leaves.  Most kernels of the late, small-spatial
    encoder blocks and of the decoder fill a fraction of the 148 SMs, so the leaves run concurrently with the chain
    instead of in line with it (measured: 33.8 → 31.1 ms per RT-1 step).  Ordering:
      * a leaf waits (event) for the main-stream kernel that produced its inputs; tensors it reads are ``record_stream``-ed
        so the caching allocator does not hand their memory to later main-stream work while the leaf may still run;
      * the "gradients ready" notification of a group of leaves (it can launch the fused reduce+AdamW of a bucket) is
        issued one group later, after the main stream has waited for the lane's event of that group — so the last leaf of
        a block still overlaps the first kernels of the next one;
      * the lane is drained (main stream waits, notifications sent) by a callback at the end of the autograd pass, so
        gradients are complete when ``backward()`` returns.
    ``RT1B200_WGRAD_STREAM=0`` runs everything in line."""

    def __init__(self, device, main):
        import os
        self.enabled = os.environ.get("RT1B200_WGRAD_STREAM", "1") != "0"
        self.device, self.main = device, main
        self.stream = torch.cuda.Stream(device=device) if self.enabled else None
        self.pending = None
        self._cb = False

    def run(self, fn, *tensors):
        if not self.enabled:
            fn()
            return
        ev = torch.cuda.Event()
        ev.record(self.main)
        self.stream.wait_event(ev)
        for t in tensors:
            if t is not None:
                t.record_stream(self.stream)
        with torch.cuda.stream(self.stream):
            fn()

    def block_done(self, params):
        """End of a group of leaves: release the previous group's notification, queue this one's."""
        if not self.enabled:
            _ready(params)
            return
        self.flush()
        ev = torch.cuda.Event()
        ev.record(self.stream)
        self.pending = (ev, params)
        if not self._cb:
            try:
                torch.autograd.Variable._execution_engine.queue_callback(self._final_flush)
                self._cb = True
            except Exception:        # not inside a backward pass
                self.flush()

    def flush(self):
        if self.enabled and self.pending is not None:
            ev, params = self.pending
            self.main.wait_event(ev)
            self.pending = None
            _ready(params)

    def _final_flush(self):
        self._cb = False
        self.flush()


_LANES = {}
import os as _os
_LINEAR_LANE = _os.environ.get("RT1B200_WGRAD_STREAM_LINEAR", "0") == "1"   # decoder linears: measured 3 % slower with the lane, off by default


def wgrad_lane(device) -> WgradLane:
    """The lane that belongs to the current stream of ``device``."""
    device = torch.device(device)
    main = torch.cuda.current_stream(device)
    key = (device.index, main.cuda_stream)
    lane = _LANES.get(key)
    if lane is None:
        lane = _LANES[key] = WgradLane(device, main)
    return lane


def _c(t: Optional[torch.Tensor]) -> Optional[torch.Tensor]:
    return None if t is None else (t if t.is_contiguous() else t.contiguous())


# ------------------------------------------------------------------------------------------------- GEMM
def gemm(a: torch.Tensor, b: torch.Tensor, out: torch.Tensor, M: int, N: int, K: int, a_mn: bool = False,
         b_mn: bool = False, bias=None, residual=None, alpha: float = 1.0, dropout_p: float = 0.0, seed: int = 0,
         split_k: int = 1, accumulate: bool = False, col_sum=None, col_sumsq=None):
    """out[M,N] = alpha * op(a) @ op(b)^T ...; operands are 2-D bf16 with unit inner stride."""
    assert a.dtype == BF16 and b.dtype == BF16 and a.stride(-1) == 1 and b.stride(-1) == 1 and out.stride(-1) == 1
    call("rt1_gemm_bf16", ptr(a), i64(a.stride(0)), i32(a_mn), ptr(b), i64(b.stride(0)), i32(b_mn), ptr(out),
         i64(out.stride(0)), i32(out.dtype == torch.float32), i32(M), i32(N), i32(K), ptr(bias), ptr(residual),
         i64(residual.stride(0) if residual is not None else 0), f32(alpha), f32(dropout_p), ctypes_u32(seed),
         i32(split_k), i32(accumulate), ptr(col_sum), ptr(col_sumsq), stream())
    return out


def ctypes_u32(v):
    import ctypes
    return ctypes.c_uint(int(v) & 0xFFFFFFFF)


class _Linear(Function):
    @staticmethod
    def forward(ctx, x, w, b, residual, dropout_p, seed, wgrad, bgrad, notify, *anchors):
        x = _c(x)
        M, K = x.shape
        N = w.shape[0]
        out = torch.empty(M, N, dtype=BF16, device=x.device)
        gemm(x, w, out, M, N, K, bias=b, residual=_c(residual), dropout_p=dropout_p, seed=seed)
        ctx.save_for_backward(x, w)
        ctx.meta = (dropout_p, seed, wgrad, bgrad, notify, b is not None, residual is not None)
        return out

    @staticmethod
    def backward(ctx, dy):
        x, w = ctx.saved_tensors
        dropout_p, seed, wgrad, bgrad, notify, has_bias, has_res = ctx.meta
        dy = _c(dy)
        M, K = x.shape
        N = w.shape[0]
        g = dy
        if dropout_p > 0.0:
            g = torch.empty_like(dy)
            call("rt1_dropout_apply", ptr(dy), ptr(g), i64(dy.numel()), f32(dropout_p), ctypes_u32(seed), stream())
        dx = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty(M, K, dtype=BF16, device=x.device)
            gemm(g, w, dx, M, K, N, a_mn=False, b_mn=True)            # dX = G · W   (W read as an [N_red, K] MN-major operand)
        def leaves():
            if wgrad is not None:
                gemm(g, x, wgrad, N, K, M, a_mn=True, b_mn=True)      # dW = Gᵀ · X  (both operands MN-major views)
            if has_bias and bgrad is not None:
                call("rt1_colsum_bf16", ptr(g), i64(g.stride(0)), i32(M), i32(N), ptr(bgrad), stream())
        if (wgrad is not None or (has_bias and bgrad is not None)) and g.is_cuda and _LINEAR_LANE:
            lane = wgrad_lane(g.device)
            lane.run(leaves, g, x)
            lane.block_done(list(notify) if notify else [])
        else:
            leaves()
            _ready(notify)
        return (dx, None, None, dy if has_res else None, None, None, None, None, None) + (None,) * (len(ctx.needs_input_grad) - 9)


def linear(x, w, b=None, residual=None, act=None, dropout_p: float = 0.0, wgrad=None, bgrad=None, notify=None):
    """``residual + dropout(x @ w.T + b)`` on the tcgen05 GEMM.  ``w``/``b`` may be ``nn.Parameter`` (their
    ``.grad`` slots receive the gradients) or plain views with explicit ``wgrad``/``bgrad`` destinations."""
    assert act is None, "activation epilogues are not used by RT-1's linears"
    anchors = []
    if wgrad is None and isinstance(w, torch.nn.Parameter) and w.requires_grad:
        wgrad = _grad_slot(w)
        anchors.append(w)
        notify = (notify or []) + [w]
    if b is not None and bgrad is None and isinstance(b, torch.nn.Parameter) and b.requires_grad:
        bgrad = _grad_slot(b)
        anchors.append(b)
        notify = (notify or []) + [b]
    if notify:
        anchors = list({id(p): p for p in list(anchors) + list(notify)}.values())
    if not torch.is_grad_enabled():
        wgrad = bgrad = None
    seed = next_seed() if dropout_p > 0 else 0
    return _Linear.apply(x, w.detach(), None if b is None else b.detach(), residual, float(dropout_p), seed, wgrad,
                         bgrad, notify, *anchors)


# -------------------------------------------------------------------------------------------- LayerNorm
class _LayerNorm(Function):
    @staticmethod
    def forward(ctx, x, w, b, eps, wp, bp):
        shp = x.shape
        D = shp[-1]
        x2 = _c(x).view(-1, D)
        rows = x2.shape[0]
        y = torch.empty_like(x2)
        mean = torch.empty(rows, dtype=torch.float32, device=x.device)
        rstd = torch.empty_like(mean)
        call("rt1_layernorm_fwd", ptr(x2), ptr(w), ptr(b), ptr(y), ptr(mean), ptr(rstd), i32(rows), i32(D), f32(eps),
             stream())
        ctx.save_for_backward(x2, w, mean, rstd)
        ctx.params = (wp, bp)
        return y.view(shp)

    @staticmethod
    def backward(ctx, dy):
        x2, w, mean, rstd = ctx.saved_tensors
        wp, bp = ctx.params
        rows, D = x2.shape
        dy2 = _c(dy).view(rows, D)
        dx = torch.empty_like(x2)
        n_scratch = min(148, (rows + 7) // 8) * 2 * D        # == rt1_layernorm_bwd_scratch_floats(rows, D)
        scratch = torch.empty(n_scratch, dtype=torch.float32, device=x2.device)
        dw, db = _grad_slot(wp), _grad_slot(bp)
        call("rt1_layernorm_bwd", ptr(x2), ptr(w), ptr(dy2), ptr(mean), ptr(rstd), ptr(dx), ptr(dw), ptr(db),
             ptr(scratch), i32(rows), i32(D), stream())
        _ready([wp, bp])
        return dx.view(dy.shape), None, None, None, None, None


def layer_norm(x, w, b, eps: float = 1e-5):
    return _LayerNorm.apply(x, w.detach(), b.detach(), float(eps), w, b)


# --------------------------------------------------------------------------------- action-token CE loss
class _ActionCE(Function):
    @staticmethod
    def forward(ctx, logits, labels, num_items):
        R, A, V = logits.shape
        logits = _c(logits)
        labels = _c(labels.to(torch.int64))
        loss = torch.empty(R, dtype=torch.float32, device=logits.device)
        row_ce = torch.empty(R * A, dtype=torch.float32, device=logits.device)
        row_lse = torch.empty_like(row_ce)
        pred = torch.empty(R, A, dtype=torch.int64, device=logits.device)
        call("rt1_action_ce_fwd", ptr(logits), ptr(labels), ptr(loss), ptr(row_ce), ptr(row_lse), ptr(pred), i32(R),
             i32(A), i32(V), f32(num_items), stream())
        ctx.save_for_backward(logits, labels, row_lse)
        ctx.num_items = num_items
        ctx.mark_non_differentiable(pred)
        return loss, pred

    @staticmethod
    def backward(ctx, dloss, _dpred):
        logits, labels, row_lse = ctx.saved_tensors
        R, A, V = logits.shape
        dlogits = torch.empty_like(logits)
        call("rt1_action_ce_bwd", ptr(logits), ptr(labels), ptr(row_lse), ptr(_c(dloss.float())), ptr(dlogits), i32(R),
             i32(A), i32(V), f32(ctx.num_items), stream())
        return dlogits, None, None


def action_ce_loss(logits, labels, num_items: float):
    return _ActionCE.apply(logits, labels, float(num_items))


# ----------------------------------------------------------------------------------------- pre-processing
def preprocess_nhwc(image: torch.Tensor, shift, ratio: float = 0.07) -> torch.Tensor:
    """``shift`` is ``(dh, dw)`` as Python ints, or an int32 device tensor ``[2]`` read by the kernel when it runs."""
    N, C, H, W = image.shape
    assert C == 3
    if image.dtype not in (torch.uint8, torch.float32):
        image = image.float()
    image = _c(image)
    out = torch.empty(N, H, W, 8, dtype=BF16, device=image.device)
    if isinstance(shift, torch.Tensor):
        assert shift.is_cuda and shift.dtype == torch.int32 and shift.numel() == 2
        call("rt1_preprocess_nhwc_dev", ptr(image), i32(image.dtype == torch.uint8), ptr(out), i32(N), i32(H), i32(W),
             ptr(shift), i32(int(H * ratio)), i32(int(W * ratio)), stream())
    else:
        call("rt1_preprocess_nhwc", ptr(image), i32(image.dtype == torch.uint8), ptr(out), i32(N), i32(H), i32(W),
             i32(shift[0]), i32(shift[1]), i32(int(H * ratio)), i32(int(W * ratio)), stream())
    return out


_seed_epoch = {}


def seed_epoch_tensor(device) -> torch.Tensor:
    """The device word added to every dropout seed on ``device`` (created and registered with the library on first use).
    ``t.add_(odd constant)`` inside a captured step gives each replay its own dropout masks."""
    device = torch.device(device)
    t = _seed_epoch.get(device.index)
    if t is None:
        t = torch.zeros(1, dtype=torch.int32, device=device)
        with torch.cuda.device(device):
            call("rt1_set_seed_epoch", ptr(t))
        _seed_epoch[device.index] = t
    return t


# ---------------------------------------------------------------------------------------------- optimizer
def fused_adamw(param_lp, master, grad, exp_avg, exp_avg_sq, h: dict, grad_scale: float = 1.0, zero_grad: bool = True):
    call("rt1_fused_adamw", ptr(param_lp), ptr(master), ptr(grad), i32(grad.dtype == BF16), ptr(exp_avg),
         ptr(exp_avg_sq), i64(master.numel()), f32(h["lr"]), f32(h["beta1"]), f32(h["beta2"]), f32(h["eps"]),
         f32(h["weight_decay"]), f32(h["bc1"]), f32(h["bc2"]), f32(grad_scale), i32(zero_grad), stream())


def l2_flush(buf: torch.Tensor, tag: int = 0):
    call("rt1_l2_flush", ptr(buf), i64(buf.numel() * buf.element_size()), ctypes_u32(tag), stream())


# ---------------------------------------------------------------- ops whose kernels land in later files
from .kernels_attn import rt1_attention, token_learner  # noqa: E402,F401
