"""NHWC sm_100a execution of the FiLM-EfficientNet-B3 encoder: one ``autograd.Function`` for the whole network.

Forward and backward are explicit kernel schedules over the kernels in ``csrc/cnn_*.cu`` and the tcgen05 GEMM
(``csrc/gemm_tcgen05.cu``); nothing here is a PyTorch op except allocation.  Per MBConv block:

  forward   expand GEMM (+BN1 statistics and their finalisation from the epilogue / last CTA) → depthwise (BN1+SiLU applied
            to the shared-memory tile, BN2 statistics + finalisation) → SE squeeze → SE FCs → SE excite (writes the GEMM
            operand) → project GEMM (+BN3 statistics + finalisation) → BN3 + stochastic depth + residual + FiLM
            (7 launches per block; evaluation mode reads cached scale / shift instead of statistics)
  backward  FiLM/BN3 backward (reduce, combine, apply) → project data gradient → one pass for the SE-scale gradient and the
            separable parts of BN2's backward sums → SE FC backward (two launches: data + four parameter gradients + BN2's
            sums) → BN2 apply → depthwise data gradient (× SiLU', BN1 sums) → BN1 apply → expand data gradient; the weight
            gradients (split-K GEMMs, depthwise wgrad) are leaves on a second stream and are written straight into the
            parameters' ``.grad`` slots.

Raw (pre-BatchNorm) tensors are what is stored for the backward pass; activated tensors are recomputed on load.
"""
from __future__ import annotations

import ctypes
from typing import List, Optional

import os

import torch
from torch.autograd import Function

from . import _lib
from ._lib import call, f32, i32, i64, ptr, stream
from .kernels import BF16, _grad_slot, _ready, gemm, wgrad_lane
from ..utils import trace

F32 = torch.float32
# BatchNorm+SiLU of the expand conv applied inside the depthwise tile (1) or materialised by an elementwise pass (0)
_DW_ACT_ON_LOAD = os.environ.get("RT1B200_DW_ACT_ON_LOAD", "1") != "0"


class _ZeroArena:
    """One pre-zeroed fp32 buffer per encoder pass from which the ~200 small statistics / accumulator tensors — and the
    split-K weight-gradient accumulators of the backward — are carved: a single fill launch instead of one per tensor
    (they were 184 launches and 0.4 ms of a 35 ms step).  Each pass gets its own arena and the views keep it alive, so
    buffers saved for the backward never alias a later pass.  The size is learnt: a pass that outgrows its arena falls
    back to separate allocations and records what it needed under ``key``; the next pass of that kind asks for that much."""
    NEED = {}

    def __init__(self, device, key=None, numel: int = 4 << 20):
        self.key = (key, torch.device(device).index)
        numel = max(numel, self.NEED.get(self.key, 0))
        self.buf = torch.zeros(numel, dtype=F32, device=device)
        self.off = 0
        self.spilled = 0

    def take(self, shape):
        n = 1
        for d in shape:
            n *= int(d)
        n_al = (n + 63) // 64 * 64
        if self.off + n_al > self.buf.numel():
            self.spilled += n_al
            return torch.zeros(*shape, dtype=F32, device=self.buf.device)
        out = self.buf[self.off:self.off + n].view(*shape)
        self.off += n_al
        return out

    def close(self):
        if self.spilled:
            self.NEED[self.key] = max(self.NEED.get(self.key, 0), self.off + self.spilled)


_ARENA = None       # set for the duration of _EncoderFn.forward / backward
_STATS_EPOCH = 0    # training passes of the encoder being run (set by _EncoderFn.forward; keys the evaluation-mode BN cache)


def _zeros(*shape, device):
    if _ARENA is not None and _ARENA.buf.device == torch.device(device):
        return _ARENA.take(shape)
    return torch.zeros(*shape, dtype=F32, device=device)


def _empty(*shape, device, dtype=BF16):
    return torch.empty(*shape, dtype=dtype, device=device)


class BnTail(ctypes.Structure):
    """``Rt1BnTail`` of include/rt1b200.h: what the statistics-producing kernel needs to finalise the BatchNorm itself."""
    _fields_ = [("gamma", ctypes.c_void_p), ("beta", ctypes.c_void_p), ("running_mean", ctypes.c_void_p),
                ("running_var", ctypes.c_void_p), ("a", ctypes.c_void_p), ("b", ctypes.c_void_p), ("mean", ctypes.c_void_p),
                ("rstd", ctypes.c_void_p), ("ticket", ctypes.c_void_p), ("count", ctypes.c_float), ("momentum", ctypes.c_float),
                ("eps", ctypes.c_float)]


# BatchNorm finalisation by the last CTA of the producing kernel (1) or by a launch of its own (0, the A/B baseline)
_BN_TAIL = os.environ.get("RT1B200_BN_TAIL", "1") != "0"


class _BN:
    """Per-layer BatchNorm state of one forward pass: batch statistics → scale/shift, kept for the backward."""
    __slots__ = ("mod", "a", "b", "mean", "rstd", "sum", "sumsq", "ticket")

    def __init__(self, mod, device):
        C = mod.num_features
        self.mod = mod
        buf = _zeros(6 * C + 16, device=device)                  # the arena hands out zeros: sums and ticket start at 0
        self.sum, self.sumsq, self.a, self.b, self.mean, self.rstd = buf[:6 * C].view(6, C).unbind(0)
        self.ticket = buf[6 * C:]

    def tail(self, count: int):
        """By-reference argument for the ``*_bn`` entry points (training mode only), or None when tails are switched off."""
        if not _BN_TAIL:
            return None
        m = self.mod
        t = BnTail(m.weight.data_ptr(), m.bias.data_ptr(), m.running_mean.data_ptr(), m.running_var.data_ptr(), self.a.data_ptr(),
                   self.b.data_ptr(), self.mean.data_ptr(), self.rstd.data_ptr(), self.ticket.data_ptr(), float(count),
                   float(m.momentum if m.momentum is not None else 0.1), float(m.eps))
        return t

    def finalize(self, count: int, training: bool, stats_epoch: int = 0):
        m = self.mod
        if not training:
            # Evaluation reads the running statistics: scale / shift only change when the statistics or the affine
            # parameters do, so they are computed once and kept on the module (78 tiny launches per policy step otherwise).
            # `stats_epoch` counts the training passes of the owning encoder — the kernels update the running statistics
            # through raw pointers, which tensor version counters do not see.
            key = (stats_epoch, m.running_mean._version, m.running_var._version, m.weight._version, m.bias._version,
                   m.weight.data_ptr(), m.running_mean.data_ptr())
            cached = getattr(m, "_rt1_eval_affine", None)
            if cached is None or cached[0] != key:
                buf = torch.empty(4, m.num_features, dtype=F32, device=m.running_mean.device)
                call("rt1_bn_finalize", ptr(None), ptr(None), f32(count), ptr(m.weight), ptr(m.bias), ptr(m.running_mean),
                     ptr(m.running_var), f32(0.0), f32(m.eps), i32(0), ptr(buf[0]), ptr(buf[1]), ptr(buf[2]), ptr(buf[3]),
                     i32(m.num_features), stream())
                cached = (key, buf)
                if not torch.cuda.is_current_stream_capturing():     # memory allocated under capture belongs to that graph
                    m._rt1_eval_affine = cached
            self.a, self.b, self.mean, self.rstd = cached[1].unbind(0)
            return
        call("rt1_bn_finalize", ptr(self.sum), ptr(self.sumsq), f32(count), ptr(m.weight), ptr(m.bias), ptr(m.running_mean),
             ptr(m.running_var), f32(m.momentum if m.momentum is not None else 0.1), f32(m.eps), i32(training), ptr(self.a),
             ptr(self.b), ptr(self.mean), ptr(self.rstd), i32(m.num_features), stream())


def _conv_bn(x2d, w2d, out2d, M, N, K, bn: "_BN", training: bool):
    """1x1 convolution whose epilogue takes the BatchNorm statistics of its output and (training) finalises them."""
    tail = bn.tail(M) if training else None
    if tail is not None:
        call("rt1_gemm_bf16_bn", ptr(x2d), i64(x2d.stride(0)), i32(0), ptr(w2d), i64(w2d.stride(0)), i32(0), ptr(out2d),
             i64(out2d.stride(0)), i32(M), i32(N), i32(K), ptr(bn.sum), ptr(bn.sumsq), ctypes.byref(tail), stream())
    else:
        gemm(x2d, w2d, out2d, M, N, K, col_sum=bn.sum if training else None, col_sumsq=bn.sumsq if training else None)
        bn.finalize(M, training, _STATS_EPOCH)


def _conv_w2d(conv) -> torch.Tensor:
    w = conv.weight
    return w.detach().view(w.shape[0], -1)


def _wgrad_gemm(g: torch.Tensor, x: torch.Tensor, param: torch.nn.Parameter, n_out: int, k_out: int, m_red: int):
    """param.grad[n_out, k_out] = g[m_red, n_out]^T @ x[m_red, k_out] — both operands read as MN-major views.
    Short outputs with a long reduction (1x1 convs over ~1M pixels) use split-K into an fp32 scratch."""
    slot = _grad_slot(param)
    tiles = ((n_out + 127) // 128) * ((k_out + 63) // 64)
    kb = (m_red + 63) // 64
    split = 1
    if tiles < 148 and kb >= 16:
        split = max(1, min(kb // 4, (148 * 2 + tiles - 1) // tiles))
    if split == 1:
        gemm(g, x, slot.view(n_out, k_out), n_out, k_out, m_red, a_mn=True, b_mn=True)
    else:
        acc = _zeros(n_out, k_out, device=g.device)
        gemm(g, x, acc, n_out, k_out, m_red, a_mn=True, b_mn=True, split_k=split)
        call("rt1_cast_f32_bf16", ptr(acc), ptr(slot), i64(acc.numel()), stream())


def _film_operands(enc):
    """Concatenated FiLM projection weights/biases [2*sum(C), 512] / [2*sum(C)] (+ their gradient destinations)."""
    w = getattr(enc, "_film_w", None)
    films = enc.film_layers()
    first = films[0]._projection_add.weight
    if w is not None and w.data_ptr() == first.data_ptr():
        return w, enc._film_b, enc._film_w_grad, enc._film_b_grad, True
    ws, bs = [], []
    for f in films:
        ws += [f._projection_add.weight, f._projection_mult.weight]
        bs += [f._projection_add.bias, f._projection_mult.bias]
    return torch.cat([t.detach() for t in ws], 0), torch.cat([t.detach() for t in bs], 0), None, None, False


class _EncoderFn(Function):
    @staticmethod
    def forward(ctx, enc, x8, context, sd_scales, *anchors):
        global _ARENA, _STATS_EPOCH
        net = enc.net
        dev = x8.device
        _ARENA = _ZeroArena(dev, ("fwd", tuple(x8.shape)))
        training = enc.training
        _STATS_EPOCH = getattr(enc, "_bn_stats_epoch", 0)
        N, H, W, _ = x8.shape
        S = []                                    # saved per-layer state for backward
        ctx.enc, ctx.training = enc, training
        context = context.contiguous()

        # ---- FiLM gamma/beta of every layer in one GEMM ----------------------------------------------
        films = enc.film_layers()
        film_off, off = [], 0
        for f in films:
            film_off.append(off)
            off += 2 * f._projection_add.out_features
        GB = off
        fw, fb, _, _, _ = _film_operands(enc)
        gb = _empty(N, GB, device=dev)
        gemm(context, fw, gb, N, GB, context.shape[1], bias=fb)
        use_early = bool(enc.early_film and net.include_film)

        # ---- stem --------------------------------------------------------------------------------------
        conv0, bn0m = net.convNormAct0[0], net.convNormAct0[1]
        C0 = conv0.out_channels
        Ho, Wo = (H + 2 - 3) // 2 + 1, (W + 2 - 3) // 2 + 1
        stem_raw = _empty(N, Ho, Wo, C0, device=dev)
        bn0 = _BN(bn0m, dev)
        # 3x3 s2 stem as im2col (K = 27 padded to 32) + tcgen05 GEMM with the BatchNorm statistics in its epilogue
        stem_cols = _empty(N * Ho * Wo, 32, device=dev)
        call("rt1_stem_im2col", ptr(x8), ptr(stem_cols), i32(N), i32(H), i32(W), stream())
        w32 = torch.nn.functional.pad(conv0.weight.detach().view(C0, 27), (0, 5))
        _conv_bn(stem_cols, w32, stem_raw.view(N * Ho * Wo, C0), N * Ho * Wo, C0, 32, bn0, training)
        if not training:
            stem_cols = None
        x = _empty(N, Ho, Wo, C0, device=dev)
        call("rt1_bn_act_scale", ptr(stem_raw), ptr(bn0.a), ptr(bn0.b), ptr(None), ptr(x), i32(N), i32(Ho * Wo), i32(C0), i32(1),
             stream())
        ctx.stem = (x8, stem_raw, bn0, stem_cols)
        Hc, Wc = Ho, Wo

        # ---- MBConv blocks -------------------------------------------------------------------------------
        for bi, blk in enumerate(net.blocks):
            if trace.ENABLED:
                torch.cuda.nvtx.range_push(f"encoder/block {bi} fwd")
            st = {}
            Cin, Ce, Co, k, s = blk.in_size, blk.expand_size, blk.out_size, blk.kernel_size, blk.strides
            M = N * Hc * Wc
            st["x_in"], st["Hin"], st["Win"] = x, Hc, Wc
            if blk.expand is not None:
                e_raw = _empty(N, Hc, Wc, Ce, device=dev)
                bn1 = _BN(blk.expand[1], dev)
                _conv_bn(x.view(M, Cin), _conv_w2d(blk.expand[0]), e_raw.view(M, Ce), M, Ce, Cin, bn1, training)
                if _DW_ACT_ON_LOAD:
                    # SiLU(BN1(.)) is applied to the depthwise kernel's shared-memory tile (once per element and CTA):
                    # the activated expand tensor never exists in HBM
                    dw_in, a_in, b_in = e_raw, bn1.a, bn1.b
                    e_act = None
                else:
                    e_act = _empty(N, Hc, Wc, Ce, device=dev)
                    call("rt1_bn_act_scale", ptr(e_raw), ptr(bn1.a), ptr(bn1.b), ptr(None), ptr(e_act), i32(N), i32(Hc * Wc),
                         i32(Ce), i32(1), stream())
                    dw_in, a_in, b_in = e_act, None, None
                st["e_raw"], st["e_act"], st["bn1"] = e_raw, e_act, bn1
            else:
                dw_in, a_in, b_in = x, None, None
            pad = (k - 1) // 2
            Ho, Wo = (Hc + 2 * pad - k) // s + 1, (Wc + 2 * pad - k) // s + 1
            HWo, Mo = Ho * Wo, N * Ho * Wo
            d_raw = _empty(N, Ho, Wo, Ce, device=dev)
            bn2 = _BN(blk.depthwise[1], dev)
            tail2 = bn2.tail(Mo) if training else None
            if tail2 is not None:
                call("rt1_dwconv_fwd_bn", ptr(dw_in), ptr(a_in), ptr(b_in), ptr(blk.depthwise[0].weight), ptr(d_raw), ptr(bn2.sum),
                     ptr(bn2.sumsq), ctypes.byref(tail2), i32(N), i32(Hc), i32(Wc), i32(Ce), i32(k), i32(s), stream())
            else:
                call("rt1_dwconv_fwd", ptr(dw_in), ptr(a_in), ptr(b_in), ptr(blk.depthwise[0].weight), ptr(d_raw),
                     ptr(bn2.sum if training else None), ptr(bn2.sumsq if training else None), i32(N), i32(Hc), i32(Wc), i32(Ce),
                     i32(k), i32(s), stream())
                bn2.finalize(Mo, training, _STATS_EPOCH)
            se = blk.se
            if se is not None:
                Cs = se.fc1.out_channels
                pooled = _zeros(N, Ce, device=dev)
                call("rt1_bn_act_pool", ptr(d_raw), ptr(bn2.a), ptr(bn2.b), ptr(pooled), i32(N), i32(HWo), i32(Ce), stream())
                z1, sc = _empty(N, Cs, device=dev, dtype=F32), _empty(N, Ce, device=dev, dtype=F32)
                call("rt1_se_fc_fwd", ptr(pooled), f32(1.0 / HWo), ptr(se.fc1.weight), ptr(se.fc1.bias), ptr(se.fc2.weight),
                     ptr(se.fc2.bias), ptr(z1), ptr(sc), i32(N), i32(Ce), i32(Cs), stream())
                st["pooled"], st["z1"], st["s"] = pooled, z1, sc
            else:
                sc = None
            a2 = _empty(N, Ho, Wo, Ce, device=dev)
            call("rt1_bn_act_scale", ptr(d_raw), ptr(bn2.a), ptr(bn2.b), ptr(sc), ptr(a2), i32(N), i32(HWo), i32(Ce), i32(1), stream())
            y_raw = _empty(N, Ho, Wo, Co, device=dev)
            bn3 = _BN(blk.project[1], dev)
            _conv_bn(a2.view(Mo, Ce), _conv_w2d(blk.project[0]), y_raw.view(Mo, Co), Mo, Co, Ce, bn3, training)
            sd = sd_scales[bi] if (training and blk.has_skip and blk.drop_rate > 0 and sd_scales is not None) else None
            out = _empty(N, Ho, Wo, Co, device=dev)
            call("rt1_block_finalize", ptr(y_raw), ptr(bn3.a), ptr(bn3.b), ptr(x if blk.has_skip else None), ptr(sd),
                 ptr(gb if use_early else None), i64(GB), i64(film_off[bi] if use_early else 0), ptr(out), i32(N), i32(HWo), i32(Co),
                 stream())
            st.update(d_raw=d_raw, bn2=bn2, a2=a2, y_raw=y_raw, bn3=bn3, sd=sd, Ho=Ho, Wo=Wo)
            S.append(st)
            x, Hc, Wc = out, Ho, Wo
            if trace.ENABLED:
                torch.cuda.nvtx.range_pop()

        # ---- head: 1x1 → BN → SiLU → 1x1 → FiLM ------------------------------------------------------------
        convh, bnhm = net.convNormAct1[0], net.convNormAct1[1]
        Cl, Ch = convh.in_channels, convh.out_channels
        M = N * Hc * Wc
        h_raw = _empty(M, Ch, device=dev)
        bnh = _BN(bnhm, dev)
        _conv_bn(x.view(M, Cl), _conv_w2d(convh), h_raw, M, Ch, Cl, bnh, training)
        t_act = _empty(M, Ch, device=dev)
        call("rt1_bn_act_scale", ptr(h_raw), ptr(bnh.a), ptr(bnh.b), ptr(None), ptr(t_act), i32(N), i32(Hc * Wc), i32(Ch), i32(1),
             stream())
        E = enc.conv1x1.out_channels
        f_raw = _empty(M, E, device=dev)
        gemm(t_act, _conv_w2d(enc.conv1x1), f_raw, M, E, Ch)
        feats = _empty(N, Hc, Wc, E, device=dev)
        call("rt1_block_finalize", ptr(f_raw), ptr(None), ptr(None), ptr(None), ptr(None), ptr(gb), i64(GB), i64(film_off[-1]),
             ptr(feats), i32(N), i32(Hc * Wc), i32(E), stream())
        ctx.head = (x, h_raw, bnh, t_act, f_raw, Hc, Wc)
        ctx.S, ctx.gb, ctx.film_off, ctx.GB, ctx.use_early, ctx.context = S, gb, film_off, GB, use_early, context
        if training:
            enc._nbt_pending = getattr(enc, "_nbt_pending", 0) + 1
            enc._bn_stats_epoch = getattr(enc, "_bn_stats_epoch", 0) + 1
        _ARENA.close()
        _ARENA = None
        return feats

    @staticmethod
    def backward(ctx, dfeats):
        global _ARENA
        enc, S, gb, film_off, GB = ctx.enc, ctx.S, ctx.gb, ctx.film_off, ctx.GB
        net = enc.net
        dev = dfeats.device
        _ARENA = _ZeroArena(dev, ("bwd", tuple(dfeats.shape)))
        lane = wgrad_lane(dev)
        dfeats = dfeats.contiguous()
        N = dfeats.shape[0]
        dgb = torch.zeros_like(gb)
        x_head, h_raw, bnh, t_act, f_raw, Hc, Wc = ctx.head
        M, E = f_raw.shape
        HW = Hc * Wc
        Ch, Cl = h_raw.shape[1], x_head.shape[-1]
        cbuf = _zeros(2, max(Ch, max(st["d_raw"].shape[-1] for st in S)), device=dev)   # c1/c2 constants
        c1, c2 = cbuf[0], cbuf[1]

        # ---- head ------------------------------------------------------------------------------------------
        Sbuf = _empty(3 * N * E, device=dev, dtype=F32)
        df_raw = _empty(M, E, device=dev)
        call("rt1_block_finalize_bwd", ptr(dfeats), ptr(f_raw), ptr(None), ptr(gb), i64(GB), i64(film_off[-1]), ptr(None), ptr(None),
             ptr(None), ptr(None), ptr(None), ptr(None), ptr(Sbuf), ptr(c1), ptr(c2), ptr(dgb), ptr(None), ptr(None), ptr(df_raw),
             ptr(None), i32(N), i32(HW), i32(E), stream())
        wc = _conv_w2d(enc.conv1x1)
        dt = _empty(M, Ch, device=dev)
        gemm(df_raw, wc, dt, M, Ch, E, b_mn=True)
        lane.run(lambda: _wgrad_gemm(df_raw, t_act, enc.conv1x1.weight, E, Ch, M), df_raw, t_act)
        sums = _empty(2, Ch, device=dev, dtype=F32)
        dh_raw = _empty(M, Ch, device=dev)
        bnm = bnh.mod
        call("rt1_bn_act_bwd", ptr(dt), ptr(h_raw), ptr(bnh.a), ptr(bnh.b), ptr(bnh.mean), ptr(bnh.rstd), ptr(None), ptr(None),
             ptr(sums), i32(0), ptr(_grad_slot(bnm.weight)), ptr(_grad_slot(bnm.bias)), ptr(dh_raw), i32(N), i32(HW),
             i32(Ch), i32(1), stream())
        convh = net.convNormAct1[0]
        dx = _empty(M, Cl, device=dev)
        gemm(dh_raw, _conv_w2d(convh), dx, M, Cl, Ch, b_mn=True)
        lane.run(lambda: _wgrad_gemm(dh_raw, x_head.view(M, Cl), convh.weight, Ch, Cl, M), dh_raw, x_head)
        lane.block_done([enc.conv1x1.weight, bnm.weight, bnm.bias, convh.weight])
        del dt, dh_raw, df_raw

        # ---- blocks, last to first -------------------------------------------------------------------------
        dout = dx.view(N, Hc, Wc, Cl)
        for bi in range(len(S) - 1, -1, -1):
            if trace.ENABLED:
                torch.cuda.nvtx.range_push(f"encoder/block {bi} bwd")
            st, blk = S[bi], net.blocks[bi]
            Cin, Ce, Co, k, s = blk.in_size, blk.expand_size, blk.out_size, blk.kernel_size, blk.strides
            Ho, Wo, Hin, Win = st["Ho"], st["Wo"], st["Hin"], st["Win"]
            HWo, Mo, Mi = Ho * Wo, N * Ho * Wo, N * Hin * Win
            bn1, bn2, bn3 = st.get("bn1"), st["bn2"], st["bn3"]
            x_in = st["x_in"]
            ready: List[torch.nn.Parameter] = []
            # BN3 + stochastic depth + residual + FiLM
            Sbuf = _empty(3 * N * Co, device=dev, dtype=F32)
            dy_raw = _empty(Mo, Co, device=dev)
            dskip = _empty(Mo, Co, device=dev) if blk.has_skip else None
            b3 = bn3.mod
            call("rt1_block_finalize_bwd", ptr(dout), ptr(st["y_raw"]), ptr(x_in if blk.has_skip else None),
                 ptr(gb if ctx.use_early else None), i64(GB), i64(film_off[bi] if ctx.use_early else 0), ptr(st["sd"]), ptr(bn3.a),
                 ptr(bn3.mean), ptr(bn3.rstd), ptr(b3.weight), ptr(b3.bias), ptr(Sbuf), ptr(c1), ptr(c2), ptr(dgb),
                 ptr(_grad_slot(b3.weight)), ptr(_grad_slot(b3.bias)), ptr(dy_raw), ptr(dskip), i32(N), i32(HWo), i32(Co), stream())
            ready += [b3.weight, b3.bias]
            # project 1x1
            wp = blk.project[0]
            da2 = _empty(Mo, Ce, device=dev)
            gemm(dy_raw, _conv_w2d(wp), da2, Mo, Ce, Co, b_mn=True)
            a2_ = st["a2"]
            lane.run(lambda: _wgrad_gemm(dy_raw, a2_.view(Mo, Ce), wp.weight, Co, Ce, Mo), dy_raw, a2_)
            ready.append(wp.weight)
            # squeeze-excite
            se = blk.se
            sc = dpool = None
            if se is not None:
                Cs = se.fc1.out_channels
                # one pass over (da2, d_raw): the SE-scale gradient ds AND the separable parts of BN2's backward sums
                parts = _empty(5, N, Ce, device=dev, dtype=F32)
                call("rt1_se_bn_bwd_reduce", ptr(da2), ptr(st["d_raw"]), ptr(bn2.a), ptr(bn2.b), ptr(bn2.mean), ptr(bn2.rstd),
                     ptr(parts), i32(N), i32(HWo), i32(Ce), stream())
                ds = parts[0]
                dpool = _empty(N, Ce, device=dev, dtype=F32)
                sums = _empty(2, Ce, device=dev, dtype=F32)
                sc = st["s"]
                # FC-pair backward, its four parameter gradients and BN2's backward sums: two launches, a CTA per 32-channel slice
                se_slots = [_grad_slot(q) for q in (se.fc1.weight, se.fc1.bias, se.fc2.weight, se.fc2.bias)]
                call("rt1_se_bwd", ptr(ds), ptr(sc), ptr(st["z1"]), ptr(st["pooled"]), f32(1.0 / HWo), ptr(se.fc1.weight),
                     ptr(se.fc2.weight), ptr(parts), ptr(_zeros(N, (Cs + 3) // 4 * 4, device=dev)), ptr(dpool), ptr(sums),
                     ptr(se_slots[0]), ptr(se_slots[1]), ptr(se_slots[2]), ptr(se_slots[3]), i32(N), i32(Ce), i32(Cs), stream())
                ready += [se.fc1.weight, se.fc1.bias, se.fc2.weight, se.fc2.bias]
            # BN2 + SiLU (+ SE scale, + pooled gradient)
            b2 = bn2.mod
            dd_raw = _empty(Mo, Ce, device=dev)
            sums_ready = 1
            if se is None:
                sums, sums_ready = _empty(2, Ce, device=dev, dtype=F32), 0
            call("rt1_bn_act_bwd", ptr(da2), ptr(st["d_raw"]), ptr(bn2.a), ptr(bn2.b), ptr(bn2.mean), ptr(bn2.rstd), ptr(sc), ptr(dpool),
                 ptr(sums), i32(sums_ready), ptr(_grad_slot(b2.weight)), ptr(_grad_slot(b2.bias)), ptr(dd_raw), i32(N),
                 i32(HWo), i32(Ce), i32(1), stream())
            ready += [b2.weight, b2.bias]
            del da2
            # depthwise conv
            dwc = blk.depthwise[0]
            has_exp = blk.expand is not None
            act_on_load = has_exp and st["e_act"] is None
            dw_in = (st["e_raw"] if act_on_load else st["e_act"]) if has_exp else x_in
            wa, wb = (bn1.a, bn1.b) if act_on_load else (None, None)
            dw_slot = _grad_slot(dwc.weight)

            def _dw_wgrad(dd_raw=dd_raw, dw_in=dw_in, dw_slot=dw_slot, k=k, s=s, Ce=Ce, Hin=Hin, Win=Win, wa=wa, wb=wb):
                scratch = _empty(k * k * Ce, device=dev, dtype=F32)
                call("rt1_dwconv_wgrad", ptr(dd_raw), ptr(dw_in), ptr(wa), ptr(wb),
                     ptr(scratch), ptr(dw_slot), i32(N), i32(Hin), i32(Win), i32(Ce), i32(k), i32(s), stream())
            lane.run(_dw_wgrad, dd_raw, dw_in)
            ready.append(dwc.weight)
            dhe = _empty(Mi, Ce, device=dev)
            if has_exp:
                sums1 = _zeros(2, Ce, device=dev)
                call("rt1_dwconv_bwd_data", ptr(dd_raw), ptr(dwc.weight), ptr(st["e_raw"]), ptr(bn1.a), ptr(bn1.b), ptr(bn1.mean),
                     ptr(bn1.rstd), ptr(dhe), ptr(sums1), i32(N), i32(Hin), i32(Win), i32(Ce), i32(k), i32(s), stream())
                b1 = bn1.mod
                de_raw = _empty(Mi, Ce, device=dev)
                call("rt1_bn_act_bwd", ptr(dhe), ptr(st["e_raw"]), ptr(bn1.a), ptr(bn1.b), ptr(bn1.mean), ptr(bn1.rstd), ptr(None),
                     ptr(None), ptr(sums1), i32(1), ptr(_grad_slot(b1.weight)), ptr(_grad_slot(b1.bias)), ptr(de_raw),
                     i32(N), i32(Hin * Win), i32(Ce), i32(0), stream())
                we = blk.expand[0]
                dxi = _empty(Mi, Cin, device=dev)
                gemm(de_raw, _conv_w2d(we), dxi, Mi, Cin, Ce, b_mn=True, residual=dskip)
                lane.run(lambda de_raw=de_raw, x_in=x_in, we=we: _wgrad_gemm(de_raw, x_in.view(Mi, Cin), we.weight, Ce, Cin, Mi),
                         de_raw, x_in)
                ready += [b1.weight, b1.bias, we.weight]
                del de_raw
            else:
                call("rt1_dwconv_bwd_data", ptr(dd_raw), ptr(dwc.weight), ptr(None), ptr(None), ptr(None), ptr(None), ptr(None),
                     ptr(dhe), ptr(None), i32(N), i32(Hin), i32(Win), i32(Ce), i32(k), i32(s), stream())
                dxi = dhe
                if dskip is not None:
                    dxi.add_(dskip.view_as(dxi))
            del dhe, dd_raw, dy_raw, dskip
            lane.block_done(ready)
            # release this block's saved activations as early as possible
            st.clear()
            dout = dxi.view(N, Hin, Win, Cin)
            if trace.ENABLED:
                torch.cuda.nvtx.range_pop()

        lane.flush()

        # ---- stem ------------------------------------------------------------------------------------------
        x8, stem_raw, bn0, stem_cols = ctx.stem
        C0 = stem_raw.shape[-1]
        Hs, Ws = stem_raw.shape[1], stem_raw.shape[2]
        b0 = bn0.mod
        sums = _empty(2, C0, device=dev, dtype=F32)
        dstem = torch.empty_like(stem_raw)
        call("rt1_bn_act_bwd", ptr(dout), ptr(stem_raw), ptr(bn0.a), ptr(bn0.b), ptr(bn0.mean), ptr(bn0.rstd), ptr(None), ptr(None),
             ptr(sums), i32(0), ptr(_grad_slot(b0.weight)), ptr(_grad_slot(b0.bias)), ptr(dstem), i32(N), i32(Hs * Ws),
             i32(C0), i32(1), stream())
        conv0 = net.convNormAct0[0]
        Ms = N * Hs * Ws
        acc = _zeros(C0, 32, device=dev)
        gemm(dstem.view(Ms, C0), stem_cols, acc, C0, 32, Ms, a_mn=True, b_mn=True, split_k=min(296, (Ms + 63) // 64))
        _grad_slot(conv0.weight).view(C0, 27).copy_(acc[:, :27])
        _ready([b0.weight, b0.bias, conv0.weight])

        # ---- FiLM projections: dW = dgb^T @ context, db = colsum(dgb) ---------------------------------------
        fw, fb, fwg, fbg, fused = _film_operands(enc)
        context = ctx.context
        films = enc.film_layers()
        if fused:
            gemm(dgb, context, fwg, GB, context.shape[1], N, a_mn=True, b_mn=True)
            call("rt1_colsum_bf16", ptr(dgb), i64(GB), i32(N), i32(GB), ptr(fbg), stream())
        else:
            wg = _empty(GB, context.shape[1], device=dev)
            bg = _empty(GB, device=dev)
            gemm(dgb, context, wg, GB, context.shape[1], N, a_mn=True, b_mn=True)
            call("rt1_colsum_bf16", ptr(dgb), i64(GB), i32(N), i32(GB), ptr(bg), stream())
            o = 0
            for f in films:
                C = f._projection_add.out_features
                for lin in (f._projection_add, f._projection_mult):
                    _grad_slot(lin.weight).copy_(wg[o:o + C])
                    _grad_slot(lin.bias).copy_(bg[o:o + C])
                    o += C
        fl = []
        for f in films:
            fl += [f._projection_add.weight, f._projection_add.bias, f._projection_mult.weight, f._projection_mult.bias]
        _ready(fl)
        ctx.S = ctx.head = ctx.stem = None
        _ARENA.close()
        _ARENA = None
        return (None, None, None, None) + (None,) * (len(ctx.needs_input_grad) - 4)


def film_efficientnet_forward(enc, image_nhwc: torch.Tensor, context: torch.Tensor) -> torch.Tensor:
    """``image_nhwc`` bf16 [N,H,W,8], ``context`` bf16 [N,512] → bf16 [N,h,w,E] through the sm_100a kernels."""
    net = enc.net
    N = image_nhwc.shape[0]
    sd = None
    if enc.training:
        # per-(block, frame) stochastic-depth scale keep/(1-p), "row" mode like torchvision's StochasticDepth
        keep = enc.__dict__.get("_sd_keep")
        if keep is None or keep.device != image_nhwc.device:      # built once (a host → device copy cannot be graph-captured)
            p = torch.tensor([b.drop_rate if b.has_skip else 0.0 for b in net.blocks], device=image_nhwc.device, dtype=F32)
            keep = enc.__dict__["_sd_keep"] = 1.0 - p
        sd = torch.bernoulli(keep[:, None].expand(-1, N)) / keep[:, None]
    anchors = [net.convNormAct0[0].weight] if torch.is_grad_enabled() and enc.training else []
    return _EncoderFn.apply(enc, image_nhwc.contiguous(), context, sd, *anchors)


def flush_bn_counters(model: torch.nn.Module):
    """``num_batches_tracked`` of the encoder's BatchNorms is advanced lazily (once per checkpoint, not 78 tiny
    kernels per step): add the pending count to every BN buffer."""
    from ..models.film_efficientnet import EfficientNetEncoder
    for mod in model.modules():
        if isinstance(mod, EfficientNetEncoder):
            pending = getattr(mod, "_nbt_pending", 0)
            if pending:
                for m in mod.modules():
                    if isinstance(m, torch.nn.modules.batchnorm._BatchNorm) and m.num_batches_tracked is not None:
                        m.num_batches_tracked += pending
                mod._nbt_pending = 0
