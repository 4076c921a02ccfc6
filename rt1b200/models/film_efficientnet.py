"""FiLM-conditioned EfficientNet-B3 image encoder.

Behavioural contract (reference, ``pytorch_robotics_transformer/film_efficientnet/``):
  * block table / rounding rules          — ``film_efficientnet_encoder.py:36-140``
  * ``SeModule``                          — ``:142-161`` (reduce dim = 0.25 * *block input* channels)
  * ``MBConvBlock``                       — ``:164-244`` (expand 1x1 → depthwise → SE → project, row-mode
                                             stochastic depth + identity skip when stride 1 and in == out)
  * ``EfficientNet`` / ``EfficientNetB3`` — ``:246-442`` (stem 3x3 s2, 26 blocks, one FiLM after every block,
                                             head 1x1 → 1536)
  * ``FilmConditioning``                  — ``film_conditioning_layer.py:23-51`` (zero-initialised projections,
                                             ``(1 + gamma) * F + beta``)
  * ``EfficientNetEncoder``               — ``pretrained_efficientnet_encoder.py:36-74``

The module tree reproduces the reference's state-dict keys exactly (``convNormAct0.{0,1}``,
``blocks.N.block.M.{0,1}``, ``blocks.N.block.K.fc{1,2}``, ``films.N._projection_{add,mult}``,
``convNormAct1.{0,1}``, ``conv1x1``, ``film_layer``), so reference checkpoints load with ``strict=True``.

Execution: ``forward`` on float32/CPU tensors is the plain-PyTorch oracle.  For CUDA bf16 tensors the
encoder switches to the NHWC sm_100a pipeline in ``rt1b200.ops.cnn`` (1x1 convs as tcgen05 GEMMs with
BN-statistics epilogues, depthwise conv fused with the BN+SiLU of its producer, SE and FiLM folded into
the neighbouring passes).
"""
from __future__ import annotations

import copy
import math
import os
from typing import List, Optional

import torch
import torch.nn as nn
import torch.nn.functional as F

from .. import ops

# (kernel, repeats, in, out, expand, stride) of the EfficientNet-B0 stage table; B3 scales it.
_B0_STAGES = (
    (3, 1, 32, 16, 1, 1),
    (3, 2, 16, 24, 6, 2),
    (5, 2, 24, 40, 6, 2),
    (3, 3, 40, 80, 6, 2),
    (5, 3, 80, 112, 6, 1),
    (5, 4, 112, 192, 6, 2),
    (3, 1, 192, 320, 6, 1),
)
# Initialiser descriptions the reference keeps from the Keras original (``film_efficientnet_encoder.py:101-117``); the PyTorch
# modules below realise them as kaiming-style fan-out normal (convolutions) and fan-out uniform (the classifier head).
CONV_KERNEL_INITIALIZER = {"class_name": "VarianceScaling", "config": {"scale": 2.0, "mode": "fan_out", "distribution": "truncated_normal"}}
DENSE_KERNEL_INITIALIZER = {"class_name": "VarianceScaling", "config": {"scale": 1.0 / 3.0, "mode": "fan_out", "distribution": "uniform"}}

DEFAULT_BLOCKS_ARGS = [
    dict(kernel_size=k, repeats=r, in_size=i, out_size=o, expand_ratio=e, id_skip=True, strides=s, se_ratio=0.25)
    for (k, r, i, o, e, s) in _B0_STAGES
]


def round_filters(filters, divisor, width_coefficient):
    """Scale a channel count by the width multiplier and snap to a multiple of ``divisor`` (never losing >10%)."""
    scaled = filters * width_coefficient
    snapped = max(divisor, int(scaled + divisor / 2) // divisor * divisor)
    if snapped < 0.9 * scaled:
        snapped += divisor
    return int(snapped)


def round_repeats(repeats, depth_coefficient):
    return int(math.ceil(depth_coefficient * repeats))


class StochasticDepthRow(nn.Module):
    """Per-sample ("row" mode) residual-branch drop with 1/(1-p) rescale; identity in eval."""

    def __init__(self, p: float):
        super().__init__()
        self.p = float(p)
        self.mode = "row"

    def forward(self, x):
        if not self.training or self.p == 0.0:
            return x
        keep = 1.0 - self.p
        noise = torch.empty([x.shape[0]] + [1] * (x.dim() - 1), dtype=x.dtype, device=x.device).bernoulli_(keep)
        if keep > 0.0:
            noise.div_(keep)
        return x * noise


class ConvNormAct(nn.Sequential):
    """conv (no bias, padding (k-1)//2) → BatchNorm2d → optional SiLU; children indexed 0/1/2 like torchvision's
    ``Conv2dNormActivation`` so the checkpoint keys line up."""

    def __init__(self, cin, cout, kernel_size=3, stride=1, groups=1, activation=True):
        layers = [nn.Conv2d(cin, cout, kernel_size, stride, (kernel_size - 1) // 2, groups=groups, bias=False),
                  nn.BatchNorm2d(cout)]
        if activation:
            layers.append(nn.SiLU(inplace=True))
        super().__init__(*layers)
        self.out_channels = cout


class SeModule(nn.Module):
    def __init__(self, expand_size, block_in_size, se_ratio=0.25):
        super().__init__()
        se_size = max(1, int(block_in_size * se_ratio))
        self.avgpool = nn.AdaptiveAvgPool2d(1)
        self.fc1 = nn.Conv2d(expand_size, se_size, 1)
        self.silu0 = nn.SiLU(inplace=True)
        self.fc2 = nn.Conv2d(se_size, expand_size, 1)
        self.act = nn.Sigmoid()

    def forward(self, inputs):
        return inputs * self.act(self.fc2(self.silu0(self.fc1(self.avgpool(inputs)))))


class MBConvBlock(nn.Module):
    def __init__(self, kernel_size, in_size, out_size, expand_ratio, id_skip, strides, se_ratio, drop_rate):
        super().__init__()
        if not (1 <= strides <= 2):
            raise ValueError("illegal stride value")
        self.kernel_size = kernel_size
        self.in_size, self.out_size = in_size, out_size
        self.expand_ratio, self.strides = expand_ratio, strides
        self.se_ratio, self.id_skip, self.drop_rate = se_ratio, id_skip, drop_rate
        expand_size = in_size * expand_ratio
        self.expand_size = expand_size
        layers: List[nn.Module] = []
        if expand_ratio != 1:
            layers.append(ConvNormAct(in_size, expand_size, 1, 1))
        layers.append(ConvNormAct(expand_size, expand_size, kernel_size, strides, groups=expand_size))
        if 0 < se_ratio <= 1:
            layers.append(SeModule(expand_size, in_size, se_ratio))
        layers.append(ConvNormAct(expand_size, out_size, 1, 1, activation=False))
        self.block = nn.Sequential(*layers)
        if drop_rate > 0:
            self.dropout = StochasticDepthRow(drop_rate)

    @property
    def has_skip(self):
        return bool(self.id_skip and self.strides == 1 and self.in_size == self.out_size)

    # named access to the stages irrespective of whether the expand conv exists
    @property
    def expand(self) -> Optional[ConvNormAct]:
        return self.block[0] if self.expand_ratio != 1 else None

    @property
    def depthwise(self) -> ConvNormAct:
        return self.block[1 if self.expand_ratio != 1 else 0]

    @property
    def se(self) -> Optional[SeModule]:
        m = self.block[2 if self.expand_ratio != 1 else 1]
        return m if isinstance(m, SeModule) else None

    @property
    def project(self) -> ConvNormAct:
        return self.block[-1]

    def forward(self, inputs):
        x = self.block(inputs)
        if self.has_skip:
            if self.drop_rate > 0:
                x = self.dropout(x)
            x = inputs + x
        return x


class FilmConditioning(nn.Module):
    def __init__(self, num_channels: int, text_vector_size: int = 512):
        super().__init__()
        self._projection_add = nn.Linear(text_vector_size, num_channels)
        self._projection_mult = nn.Linear(text_vector_size, num_channels)
        for lin in (self._projection_add, self._projection_mult):
            nn.init.zeros_(lin.weight)
            nn.init.zeros_(lin.bias)

    def forward(self, conv_filters: torch.Tensor, conditioning: torch.Tensor):
        beta = self._projection_add(conditioning)[:, :, None, None]
        gamma = self._projection_mult(conditioning)[:, :, None, None]
        return (1 + gamma) * conv_filters + beta


class EfficientNet(nn.Module):
    def __init__(self, width_coefficient, depth_coefficient, dropout_rate=0.2, drop_connect_rate=0.2,
                 depth_divisor=8, blocks_args="default", include_top=True, classes=1000,
                 include_film=False, text_vector_size=512):
        super().__init__()
        self.dropout_rate = dropout_rate
        self.include_top = include_top
        self.include_film = include_film
        if blocks_args == "default":
            blocks_args = DEFAULT_BLOCKS_ARGS
        stages = copy.deepcopy(blocks_args)

        stem_out = round_filters(32, depth_divisor, width_coefficient)
        self.convNormAct0 = ConvNormAct(3, stem_out, 3, 2)

        total = float(sum(round_repeats(s["repeats"], depth_coefficient) for s in stages))
        blocks, films, idx = [], [], 0
        for st in stages:
            assert st["repeats"] > 0
            cin = round_filters(st["in_size"], depth_divisor, width_coefficient)
            cout = round_filters(st["out_size"], depth_divisor, width_coefficient)
            for j in range(round_repeats(st["repeats"], depth_coefficient)):
                blocks.append(MBConvBlock(
                    kernel_size=st["kernel_size"], in_size=cin if j == 0 else cout, out_size=cout,
                    expand_ratio=st["expand_ratio"], id_skip=st["id_skip"],
                    strides=st["strides"] if j == 0 else 1, se_ratio=st["se_ratio"],
                    drop_rate=drop_connect_rate * idx / total))
                if include_film:
                    films.append(FilmConditioning(cout, text_vector_size))
                idx += 1
        self.blocks = nn.ModuleList(blocks)
        if include_film:
            self.films = nn.ModuleList(films)
        head_out = round_filters(1280, depth_divisor, width_coefficient)
        self.convNormAct1 = ConvNormAct(cout, head_out, 1, 1)
        if include_top:
            self.glovalAvePool = nn.AdaptiveAvgPool2d(1)
            if dropout_rate > 0:
                self.dropout = nn.Dropout(dropout_rate)
            self.fc = nn.Linear(head_out, classes)

    def forward(self, inputs, context=None):
        x = self.convNormAct0(inputs)
        if self.include_film:
            for blk, film in zip(self.blocks, self.films):
                x = film(blk(x), context)
        else:
            for blk in self.blocks:
                x = blk(x)
        x = self.convNormAct1(x)
        if not self.include_top:
            return x
        x = self.glovalAvePool(x)
        if self.dropout_rate > 0:
            x = self.dropout(x)
        return self.fc(torch.flatten(x, 1))


_CKPT_DIR = os.path.join(os.path.dirname(__file__), "efficientnet_checkpoints")


def load_official_pytorch_param(model: nn.Module, weights_path: str) -> nn.Module:
    """Load a torchvision ``efficientnet_b3`` state dict: torchvision and this module register tensors in the
    same order, so tensors are matched positionally (reference ``film_efficientnet_encoder.py:411-425``)."""
    official = torch.load(weights_path, map_location="cpu")
    own = model.state_dict()
    remapped = {dst: official[src] for src, dst in zip(list(official), list(own))}
    own.update(remapped)
    model.load_state_dict(own)
    return model


def maybe_restore_with_film(*args, weights="imagenet", include_top=True, include_film=False,
                            weights_path: Optional[str] = None, **kwargs):
    assert weights in (None, "imagenet"), "Set weights to either None or 'imagenet'."
    if weights is None:
        return EfficientNet(*args, include_top=include_top, include_film=include_film, **kwargs)
    plain = EfficientNet(*args, include_top=include_top, include_film=False, **kwargs)
    if weights_path is None:
        weights_path = os.path.join(_CKPT_DIR, "efficientnetb3.pth" if include_top else "efficientnetb3_notop.pth")
    if not os.path.exists(weights_path):
        raise FileNotFoundError(
            f"ImageNet EfficientNet-B3 weights not found at {weights_path}; pass weights=None for random init "
            "(the upstream repository does not ship this blob either).")
    load_official_pytorch_param(plain, weights_path)
    if not include_film:
        return plain
    filmed = EfficientNet(*args, include_top=include_top, include_film=True, **kwargs)
    sd = filmed.state_dict()
    sd.update({k: v for k, v in plain.state_dict().items() if k in sd})
    filmed.load_state_dict(sd)
    return filmed


def EfficientNetB3(weights="imagenet", include_top=True, classes=1000, include_film=False, **kwargs):
    return maybe_restore_with_film(1.2, 1.4, 0.3, weights=weights, include_top=include_top, classes=classes,
                                   include_film=include_film, **kwargs)


class ILSVRCPredictor:
    """Top-k ImageNet label names of a classifier output (reference ``film_efficientnet_encoder.py:445-459``).

    The reference reads ``efficientnet_checkpoints/imagenet_classes.json`` (``{"281": ["n02123045", "tabby"], ...}``).  That
    file is looked up next to this module, in the reference installation, or at ``class_index_path``; without it the category
    names that torchvision ships with its EfficientNet-B3 weight metadata are used (same order, no download involved)."""

    def __init__(self, top: int = 5, class_index_path: Optional[str] = None):
        import json
        self.top = top
        candidates = [class_index_path, os.path.join(_CKPT_DIR, "imagenet_classes.json"),
                      os.path.join(os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))), "baseline", "_ref",
                                   "pytorch_robotics_transformer", "film_efficientnet", "efficientnet_checkpoints",
                                   "imagenet_classes.json")]
        self.class_index = None
        for c in candidates:
            if c and os.path.isfile(c):
                with open(c, "r") as f:
                    self.class_index = {int(k): v[1] for k, v in json.load(f).items()}
                break
        if self.class_index is None:
            from torchvision.models import EfficientNet_B3_Weights
            self.class_index = dict(enumerate(EfficientNet_B3_Weights.IMAGENET1K_V1.meta["categories"]))

    def predict_topk(self, out: torch.Tensor):
        assert out.shape[0] == 1, "Only accept tensor with batch size 1 to simplify implementation"
        top = torch.argsort(out[0].detach().float().cpu(), descending=True)[:self.top].tolist()
        return [self.class_index[i] for i in top]


class EfficientNetEncoder(nn.Module):
    """EfficientNet-B3 (no top, FiLM after every block) → 1x1 conv 1536→``token_embedding_size`` → FiLM.

    ``weights`` defaults to ``None`` here (random init): the ImageNet blob the reference expects is not
    distributed with it (``.MISSING_LARGE_BLOBS``); pass ``weights='imagenet'`` with the file in place to
    reproduce ``pretrained_efficientnet_encoder.py:52``.
    """

    def __init__(self, token_embedding_size: int = 512, weights: Optional[str] = None, early_film: bool = True,
                 include_top: bool = False, pooling: bool = True):
        super().__init__()
        self.conv1x1 = nn.Conv2d(1536, token_embedding_size, 1, bias=False)
        self.net = EfficientNetB3(weights=weights, include_top=include_top, include_film=early_film)
        self.film_layer = FilmConditioning(token_embedding_size, 512)
        self.early_film = early_film
        self._pooling = pooling

    def film_layers(self) -> List[FilmConditioning]:
        """All FiLM layers in execution order: one per MBConv block (early FiLM), then the final one."""
        early = list(self.net.films) if (self.early_film and self.net.include_film) else []
        return early + [self.film_layer]

    def flat_groups(self):
        # every FiLM projection of the encoder is consumed by ONE GEMM on the sm_100a path:
        # [gamma|beta for all layers] = context @ _film_w.T + _film_b ; per layer the rows are [add (beta) | mult (gamma)]
        ws, bs = [], []
        for f in self.film_layers():
            ws += [f._projection_add.weight, f._projection_mult.weight]
            bs += [f._projection_add.bias, f._projection_mult.bias]
        return {"_film_w": ws, "_film_b": bs}

    def forward(self, image: torch.Tensor, context: torch.Tensor) -> torch.Tensor:
        feats = self.net(image, context) if self.early_film else self.net(image)
        feats = self.film_layer(self.conv1x1(feats), context)
        return feats if not self._pooling else feats.mean(dim=(2, 3))

    def forward_nhwc(self, image_nhwc: torch.Tensor, context: torch.Tensor) -> torch.Tensor:
        """sm_100a path: ``image_nhwc`` bf16 ``[N,H,W,8]`` (3 channels zero-padded to 8) → bf16 ``[N,h,w,E]``."""
        return ops.cnn.film_efficientnet_forward(self, image_nhwc, context)
