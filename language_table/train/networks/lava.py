"""LAVA: language-attends-to-vision sequence policy (reference ``language_table/train/networks/lava.py``, Flax) in PyTorch.

Pipeline per frame: image encoder (conv/max-pool tower, or a frozen ResNet pyramid topped by two conv/pool levels) →
"visual sentence" (selected pyramid levels projected to ``d_model``, scaled by sqrt(d_model), fixed 2-D sin/cos positions)
→ ``num_layers`` pre-norm cross-attention blocks in which ONE language token queries the visual sentence (residual on the
language path only) → LayerNorm; then over the ``sequence_length`` frames a pre-norm temporal transformer (1-D sinusoidal
positions, mean over time, LayerNorm) → dense ResNet → action.  The language token is a sentence embedding found in the
observation (``clip_embedding``, the default) or comes from the in-graph frozen CLIP text tower
(``lang_encoder="clip"``: :mod:`.clip_text` restates the architecture; its pretrained checkpoint is not available offline —
load one with ``clip_text.load_clip_text_weights`` or pass any module mapping token ids to a 512-d embedding as
``text_encoder``).
"""
import math
from typing import Optional, Sequence

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

from . import resnet_v1
from .dense_resnet import DenseResnet, dense


class ConvMaxpoolCNNEncoder(nn.Module):
    CHANNELS = (32, 64, 128, 256)

    def __init__(self):
        super().__init__()
        cs = (3,) + self.CHANNELS
        self.convs = nn.ModuleList([nn.Conv2d(cs[i], cs[i + 1], 3, padding=1) for i in range(4)])
        self.pyramid_channels = list(self.CHANNELS) + [self.CHANNELS[-1]]

    def forward(self, rgb):                        # NCHW in [0, 1]
        x, pyramid = rgb, []
        for conv in self.convs:
            x = F.max_pool2d(F.relu(conv(x)), 2, 2)
            pyramid.append(x)
        pyramid.append(F.max_pool2d(x, 2, 2))
        return pyramid


class ResNetVisualEncoder(nn.Module):
    """Frozen two-stage bottleneck ResNet (evaluated with running statistics) + two trainable conv/pool levels."""
    MEAN, STD = (0.485, 0.456, 0.406), (0.229, 0.224, 0.225)

    def __init__(self):
        super().__init__()
        self.resnet = resnet_v1.MultiscaleResNet(stage_sizes=(3, 4), block_cls=resnet_v1.BottleneckResNetBlock)
        for p in self.resnet.parameters():
            p.requires_grad_(False)
        c = self.resnet.out_channels               # 512
        self.convs = nn.ModuleList([nn.Conv2d(c, 128, 3, padding=1), nn.Conv2d(128, 256, 3, padding=1)])
        self.pyramid_channels = [64, 64, 128, 256, 256]
        self.register_buffer("mean", torch.tensor(self.MEAN).view(1, 3, 1, 1), persistent=False)
        self.register_buffer("std", torch.tensor(self.STD).view(1, 3, 1, 1), persistent=False)

    def train(self, mode: bool = True):
        super().train(mode)
        self.resnet.eval()                         # frozen: always inference-mode BatchNorm (reference :60-62)
        return self

    def forward(self, rgb):
        feats = self.resnet(normalize_image_resnet(rgb, self.mean, self.std))
        pyramid = [feats[0], feats[1]]
        x = feats[-1]
        for conv in self.convs:
            x = F.max_pool2d(F.relu(conv(x)), 2, 2)
            pyramid.append(x)
        pyramid.append(F.max_pool2d(x, 2, 2))
        return pyramid


def normalize_image_resnet(images: torch.Tensor, mean=None, std=None) -> torch.Tensor:
    """NCHW images in [0, 1] → ImageNet-normalised, bilinearly resized (no antialiasing) to 224 × 224: the input
    convention of the pretrained ResNet (reference ``lava.py:90-98``, NHWC there)."""
    if mean is None:
        mean = torch.tensor(ResNetVisualEncoder.MEAN, device=images.device, dtype=images.dtype).view(1, 3, 1, 1)
        std = torch.tensor(ResNetVisualEncoder.STD, device=images.device, dtype=images.dtype).view(1, 3, 1, 1)
    return F.interpolate((images - mean) / std, size=(224, 224), mode="bilinear", align_corners=False)


def sinusoid_1d(length: int, dim: int, max_timescale: float = 1.0e4) -> torch.Tensor:
    pos = torch.arange(length, dtype=torch.float32)[:, None]
    div = torch.exp(torch.arange(0, dim, 2, dtype=torch.float32) * -(math.log(max_timescale) / dim))
    pe = torch.zeros(length, dim)
    pe[:, 0::2], pe[:, 1::2] = torch.sin(pos * div), torch.cos(pos * div)
    return pe[None]


def sinusoidal_init(max_len: int, max_timescale: float = 1.0e4):
    """Initialiser-shaped access to ``sinusoid_1d``: ``sinusoidal_init(n)(key, shape, dtype)`` → ``[1, n, shape[-1]]``
    (the calling convention of the reference's version, ``lava.py:101-143``; ``key`` is ignored — the table is fixed)."""
    def init(key=None, shape=None, dtype=torch.float32):
        return sinusoid_1d(max_len, shape[-1], max_timescale).to(dtype or torch.float32)
    return init


class Add1DPositionEmbedding(nn.Module):
    """``[b, t, d]`` + the fixed sinusoidal code of the first ``t`` positions (reference ``lava.py:146-186``).  The table
    has ``rescale_from[0]`` rows when given, else ``max_len``, else it is built for the input's length; nothing is learned."""

    def __init__(self, rescale_from: Optional[Sequence[int]] = None, max_len: Optional[int] = None):
        super().__init__()
        self.rescale_from, self.max_len = rescale_from, max_len
        self._table = None

    def forward(self, inputs: torch.Tensor) -> torch.Tensor:
        if inputs.dim() != 3:
            raise ValueError("Number of dimensions should be 3, but it is: %d" % inputs.dim())
        length, dim = inputs.shape[1], inputs.shape[2]
        rows = self.rescale_from[0] if self.rescale_from else (self.max_len or length)
        t = self._table
        if t is None or t.shape[1] != rows or t.shape[2] != dim or t.device != inputs.device or t.dtype != inputs.dtype:
            t = self._table = sinusoidal_init(rows)(None, (1, rows, dim), None).to(device=inputs.device, dtype=inputs.dtype)
        return inputs + t[:, :length]


def positional_encoding2d(d_model: int, height: int, width: int) -> torch.Tensor:
    """Fixed 2-D sin/cos code: first half of the channels encodes x, second half y.  The reference flattens the
    ``[d_model, H, W]`` table with a plain reshape to ``[H*W, d_model]`` (``lava.py:189-218``); that layout is kept."""
    if d_model % 4:
        raise ValueError(f"sin/cos positional encoding needs d_model % 4 == 0 (got {d_model})")
    half = d_model // 2
    div = np.exp(np.arange(0.0, half, 2) * -(np.log(10000.0) / half))
    pw, ph = np.arange(0.0, width)[:, None] * div, np.arange(0.0, height)[:, None] * div      # [W, half/2], [H, half/2]
    pe = np.zeros((d_model, height, width), dtype=np.float32)
    pe[0:half:2] = np.sin(pw).T[:, None, :]
    pe[1:half:2] = np.cos(pw).T[:, None, :]
    pe[half::2] = np.sin(ph).T[:, :, None]
    pe[half + 1::2] = np.cos(ph).T[:, :, None]
    return torch.from_numpy(pe.reshape(height * width, d_model))[None]


class VisualDescriptorsNet(nn.Module):
    def __init__(self, pyramid_channels: Sequence[int], pyramid_fuse_layers: Sequence[int], d_model: int):
        super().__init__()
        self.layers, self.d_model = tuple(pyramid_fuse_layers), d_model
        self.proj = nn.ModuleList([dense(pyramid_channels[i], d_model) for i in self.layers])

    def forward(self, pyramid):
        out = []
        for proj, idx in zip(self.proj, self.layers):
            f = pyramid[idx]                                       # [n, c, h, w]
            n, c, h, w = f.shape
            x = proj(f.permute(0, 2, 3, 1).reshape(n, h * w, c)) * math.sqrt(float(self.d_model))
            out.append(x + positional_encoding2d(self.d_model, h, w).to(x))
        return torch.cat(out, dim=1)


class _FFN(nn.Module):
    def __init__(self, d_model: int, dff: int, p: float):
        super().__init__()
        self.fc1, self.fc2, self.drop = dense(d_model, dff), dense(dff, dff), nn.Dropout(p)

    def forward(self, x):
        return self.drop(self.fc2(F.relu(self.fc1(x))))


class PrenormPixelLangEncoder(nn.Module):
    """Language queries pixels; the residual runs on the language path only."""

    def __init__(self, d_model: int, num_heads: int, dropout_rate: float, mha_dropout_rate: float, dff: int):
        super().__init__()
        self.ln_pix, self.ln_lang, self.ln_ffn = nn.LayerNorm(d_model), nn.LayerNorm(d_model), nn.LayerNorm(d_model)
        self.mha = nn.MultiheadAttention(d_model, num_heads, dropout=mha_dropout_rate, batch_first=True)
        self.drop, self.ffn = nn.Dropout(dropout_rate), _FFN(d_model, dff, dropout_rate)

    def forward(self, pixel_x, lang_x):
        p, q = self.ln_pix(pixel_x), self.ln_lang(lang_x)
        x3 = lang_x + self.drop(self.mha(q, p, p, need_weights=False)[0])
        return x3 + self.ffn(self.ln_ffn(x3))


class PrenormEncoderLayer(nn.Module):
    def __init__(self, d_model: int, num_heads: int, dropout_rate: float, mha_dropout_rate: float, dff: int):
        super().__init__()
        self.ln1, self.ln2 = nn.LayerNorm(d_model), nn.LayerNorm(d_model)
        self.mha = nn.MultiheadAttention(d_model, num_heads, dropout=mha_dropout_rate, batch_first=True)
        self.drop, self.ffn = nn.Dropout(dropout_rate), _FFN(d_model, dff, dropout_rate)

    def forward(self, x):
        h = self.ln1(x)
        x3 = x + self.drop(self.mha(h, h, h, need_weights=False)[0])
        return x3 + self.ffn(self.ln2(x3))


class TemporalTransformer(nn.Module):
    def __init__(self, num_layers: int, d_model: int, num_heads: int, dff: int, sequence_length: int):
        super().__init__()
        self.d_model, self.inp, self.drop = d_model, dense(d_model, d_model), nn.Dropout(0.1)
        self.pos = Add1DPositionEmbedding(max_len=sequence_length)
        self.layers = nn.ModuleList([PrenormEncoderLayer(d_model, num_heads, 0.1, 0.0, dff) for _ in range(num_layers)])
        self.norm = nn.LayerNorm(d_model)

    def forward(self, x):                              # [b, t, d]
        x = self.inp(x) * math.sqrt(self.d_model)
        x = self.drop(self.pos(x))
        for layer in self.layers:
            x = layer(x)
        return self.norm(x.mean(dim=1))


class SequenceLAVAEncoder(nn.Module):
    def __init__(self, image_encoder: str = "conv_maxpool", lang_encoder: str = "clip_in_obs", num_layers: int = 2,
                 sequence_length: int = 4, temporal_transformer_num_layers: int = 2, d_model: int = 128, num_heads: int = 2,
                 pyramid_fuse_layers: Sequence[int] = (2, 3, 4), lang_dim: int = 512, text_encoder: Optional[nn.Module] = None):
        super().__init__()
        if image_encoder == "resnet":
            self.image_encoder = ResNetVisualEncoder()
        elif image_encoder == "conv_maxpool":
            self.image_encoder = ConvMaxpoolCNNEncoder()
        else:
            raise NotImplementedError(image_encoder)
        if lang_encoder not in ("clip_in_obs", "clip"):
            raise NotImplementedError(lang_encoder)
        if lang_encoder == "clip" and text_encoder is None:
            # the reference's in-graph frozen CLIP text tower (lava.py:423-432): same architecture, restated in clip_text.py.
            # Its pretrained checkpoint is not reachable offline: load one with clip_text.load_clip_text_weights(), or pass a
            # ready module as `text_encoder`; without either the tower is randomly initialised (and frozen).
            from language_table.train.networks.clip_text import ClipTextEncoder
            text_encoder = ClipTextEncoder(out_features=lang_dim).freeze()
        self.lang_encoder, self.text_encoder, self.d_model = lang_encoder, text_encoder, d_model
        self.descriptors = VisualDescriptorsNet(self.image_encoder.pyramid_channels, pyramid_fuse_layers, d_model)
        self.vis_drop, self.lang_drop = nn.Dropout(0.1), nn.Dropout(0.1)
        self.lang_proj = dense(lang_dim, d_model)
        self.fusion = nn.ModuleList([PrenormPixelLangEncoder(d_model, 2, 0.1, 0.0, d_model) for _ in range(num_layers)])
        self.fused_norm = nn.LayerNorm(d_model)
        self.temporal = TemporalTransformer(temporal_transformer_num_layers, d_model, num_heads, d_model, sequence_length)

    def forward(self, obs):
        rgb = obs["rgb"]                                           # [b, t, h, w, 3] in [0, 1]
        b, t, h, w, c = rgb.shape
        frames = rgb.reshape(b * t, h, w, c).permute(0, 3, 1, 2)
        sentence = self.vis_drop(self.descriptors(self.image_encoder(frames)))
        if self.lang_encoder == "clip_in_obs":
            lang = obs["clip_embedding"].reshape(b * t, -1)
        else:
            emb = self.text_encoder(obs["instruction_tokenized_clip"][:, 0].long())
            emb = emb / emb.norm(dim=-1, keepdim=True)
            lang = emb[:, None, :].expand(-1, t, -1).reshape(b * t, -1)
        q = self.lang_drop(self.lang_proj(lang) * math.sqrt(self.d_model))[:, None, :]
        for layer in self.fusion:
            q = layer(sentence, q)
        return self.temporal(self.fused_norm(q[:, 0]).reshape(b, t, -1))


class SequenceLAVMSE(nn.Module):
    def __init__(self, action_size: int, dense_resnet_width: int, dense_resnet_num_blocks: int, lava_num_layers: int,
                 lava_sequence_length: int, lava_temporal_transformer_num_layers: int, lava_d_model: int, lava_num_heads: int,
                 lava_pyramid_fuse_layers: Sequence[int], lava_image_encoder: str, lava_lang_encoder: str,
                 text_encoder: Optional[nn.Module] = None):
        super().__init__()
        self.encoder = SequenceLAVAEncoder(
            image_encoder=lava_image_encoder, lang_encoder=lava_lang_encoder, num_layers=lava_num_layers,
            sequence_length=lava_sequence_length, temporal_transformer_num_layers=lava_temporal_transformer_num_layers,
            d_model=lava_d_model, num_heads=lava_num_heads, pyramid_fuse_layers=lava_pyramid_fuse_layers, text_encoder=text_encoder)
        self.dense_resnet = DenseResnet(lava_d_model, dense_resnet_width, dense_resnet_num_blocks)
        self.action_projection = dense(dense_resnet_width, action_size)

    def forward(self, obs):
        return self.action_projection(self.dense_resnet(self.encoder(obs)))
