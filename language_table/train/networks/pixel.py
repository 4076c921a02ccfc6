"""Pixel + language MSE baseline (reference ``language_table/train/networks/pixel.py:25-111``): frames stacked
channel-wise, conv/max-pool tower with multiplicative language fusion from the second conv on, dense ResNet head."""
import torch
import torch.nn as nn
import torch.nn.functional as F

from .dense_resnet import DenseResnet, dense


class LanguageFusion(nn.Module):
    def __init__(self, lang_dim: int, channels: int):
        super().__init__()
        self.proj = dense(lang_dim, channels)

    def forward(self, lang, image):           # image NCHW
        return image * self.proj(lang)[:, :, None, None]


class ConvMaxpoolLanguageEncoder(nn.Module):
    CHANNELS = (32, 64, 128, 256)
    FUSE_FROM = 2

    def __init__(self, in_channels: int, lang_dim: int = 512):
        super().__init__()
        self.convs, self.fusions, c = nn.ModuleList(), nn.ModuleList(), in_channels
        for idx, ch in enumerate(self.CHANNELS):
            self.convs.append(nn.Conv2d(c, ch, 3, padding=1))
            self.fusions.append(LanguageFusion(lang_dim, ch) if self.FUSE_FROM <= idx + 1 else nn.Identity())
            c = ch
        self.final_lang = dense(lang_dim, self.CHANNELS[-1])
        self.norm = nn.LayerNorm(self.CHANNELS[-1])
        self.out_features = self.CHANNELS[-1]

    def forward(self, rgb, lang):
        x = rgb
        for conv, fuse in zip(self.convs, self.fusions):
            x = conv(x)
            if isinstance(fuse, LanguageFusion):
                x = fuse(lang, x)
            x = F.max_pool2d(F.relu(x), 2, 2)
        x = x.mean(dim=(2, 3)) * self.final_lang(lang)
        return self.norm(F.relu(x))


class PixelLangMSE(nn.Module):
    def __init__(self, action_size: int, dense_resnet_width: int, dense_resnet_num_blocks: int, sequence_length: int = 4,
                 lang_dim: int = 512):
        super().__init__()
        self.encoder = ConvMaxpoolLanguageEncoder(3 * sequence_length, lang_dim)
        self.dense_resnet = DenseResnet(self.encoder.out_features, dense_resnet_width, dense_resnet_num_blocks)
        self.action_projection = dense(dense_resnet_width, action_size)

    def forward(self, obs):
        rgb = obs["rgb"]                                     # [b, n, h, w, 3] in [0, 1]
        b, n, h, w, c = rgb.shape
        rgb = rgb.permute(0, 1, 4, 2, 3).reshape(b, n * c, h, w)
        lang = obs["clip_embedding"][:, -1]
        return self.action_projection(self.dense_resnet(self.encoder(rgb, lang)))
