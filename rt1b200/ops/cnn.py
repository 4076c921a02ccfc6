"""NHWC pipeline of the FiLM-EfficientNet-B3 image encoder (sm_100a path).

Entry point: :func:`film_efficientnet_forward` — called by ``EfficientNetEncoder.forward_nhwc`` with bf16
NHWC frames.  It walks the parameter containers in ``rt1b200.models.film_efficientnet`` and runs each stage
through the NHWC kernels of ``rt1b200.ops.kernels_cnn`` (K2–K8 of SURVEY §2.6):

  stem      : im2col + tcgen05 GEMM (BN statistics + their finalisation in the epilogue) → BN + SiLU pass
  MBConv    : [expand 1x1 GEMM + BN1 statistics/tail] → [depthwise tile kernel: SiLU(BN1(.)) applied to the shared-memory
              tile, BN2 statistics/tail] → [BN2+SiLU + SE pool] → [SE FCs] → [BN2+SiLU+SE-scale pass] → [project GEMM +
              BN3 statistics/tail] → [BN3 + stochastic depth + residual + FiLM affine]
  head      : 1x1 384→1536 + BN + SiLU → 1x1 1536→512 → FiLM
(the schedule itself, forward and backward, is ``rt1b200/ops/kernels_cnn.py``)

``film_efficientnet_forward_torch`` is the same network on stock channels-last PyTorch ops (cuDNN / cuBLAS bf16).  It is
NOT a product path: it exists as the library baseline of ``bench.py --cnn cudnn`` and as the bisecting arm of the tests
(``force_torch_encoder``); on a GPU the product path is always the sm_100a kernels, and a missing library raises.
"""
from __future__ import annotations

import os
from typing import Optional

import torch
import torch.nn.functional as F


def _bn(x, bn, training: bool):
    """BatchNorm on a bf16 NCHW(channels-last) tensor with fp32 running statistics."""
    w = bn.weight.float() if bn.weight is not None else None
    b = bn.bias.float() if bn.bias is not None else None
    if training and bn.track_running_stats and bn.num_batches_tracked is not None:
        bn.num_batches_tracked.add_(1)
    use_batch = training or not bn.track_running_stats
    return F.batch_norm(x, bn.running_mean if bn.track_running_stats else None,
                        bn.running_var if bn.track_running_stats else None, w, b,
                        use_batch, bn.momentum if bn.momentum is not None else 0.1, bn.eps)


def _cna(x, cna, training: bool):
    conv, bn = cna[0], cna[1]
    x = F.conv2d(x, conv.weight, None, conv.stride, conv.padding, conv.dilation, conv.groups)
    x = _bn(x, bn, training)
    if len(cna) > 2:
        x = F.silu(x)
    return x


def _film(x, film, ctx):
    beta = F.linear(ctx, film._projection_add.weight, film._projection_add.bias)
    gamma = F.linear(ctx, film._projection_mult.weight, film._projection_mult.bias)
    return (1 + gamma[:, :, None, None]) * x + beta[:, :, None, None]


def _mbconv_torch(x, blk, training: bool):
    inp = x
    if blk.expand is not None:
        x = _cna(x, blk.expand, training)
    x = _cna(x, blk.depthwise, training)
    se = blk.se
    if se is not None:
        s = x.mean(dim=(2, 3), keepdim=True)
        s = F.silu(F.conv2d(s, se.fc1.weight, se.fc1.bias))
        s = torch.sigmoid(F.conv2d(s, se.fc2.weight, se.fc2.bias))
        x = x * s
    x = _cna(x, blk.project, training)
    if blk.has_skip:
        if blk.drop_rate > 0 and training:
            x = blk.dropout(x)
        x = inp + x
    return x


def film_efficientnet_forward_torch(encoder, image_nhwc: torch.Tensor, context: torch.Tensor) -> torch.Tensor:
    """Channels-last PyTorch execution of the encoder on bf16 tensors (bring-up / bisecting path)."""
    net = encoder.net
    training = encoder.training
    x = image_nhwc[..., :3].permute(0, 3, 1, 2)                 # NCHW view of NHWC memory (channels_last)
    x = _cna(x, net.convNormAct0, training)
    for i, blk in enumerate(net.blocks):
        x = _mbconv_torch(x, blk, training)
        if net.include_film:
            x = _film(x, net.films[i], context)
    x = _cna(x, net.convNormAct1, training)
    x = F.conv2d(x, encoder.conv1x1.weight)
    x = _film(x, encoder.film_layer, context)
    return x.permute(0, 2, 3, 1).contiguous()


_FORCE_TORCH = os.environ.get("RT1B200_CNN_BENCH_CUDNN", "0") == "1"     # set by `bench.py --cnn cudnn` only


class force_torch_encoder:
    """Test/bench helper: run the encoder on stock PyTorch ops inside the ``with`` block."""

    def __enter__(self):
        global _FORCE_TORCH
        self._prev, _FORCE_TORCH = _FORCE_TORCH, True

    def __exit__(self, *exc):
        global _FORCE_TORCH
        _FORCE_TORCH = self._prev


def film_efficientnet_forward(encoder, image_nhwc: torch.Tensor, context: torch.Tensor) -> torch.Tensor:
    from . import _lib
    if _FORCE_TORCH or os.environ.get("RT1B200_CNN_BENCH_CUDNN", "0") == "1":
        return film_efficientnet_forward_torch(encoder, image_nhwc, context)
    if not _lib.available():
        raise RuntimeError(f"rt1b200: the native kernel library is not loadable ({_lib.load_error()})")
    from . import kernels_cnn
    return kernels_cnn.film_efficientnet_forward(encoder, image_nhwc, context)
