"""Image pre-processing: dtype conversion + pad-and-shift "random crop".

Behavioural contract: reference ``film_efficientnet/preprocessors.py:37-56`` — uint8 → float/255, zero-pad by
``int(ratio * size)`` on each side, draw ONE (dh, dw) offset for the whole batch, read back an H x W window.
The reference applies it in eval as well (``transformer_network.py:445``); ``TransformerNetwork`` keeps that
default and exposes ``eval_center_crop`` to make inference deterministic.

The offset can be injected (``shift=(dh, dw)``) so kernels and the oracle see identical crops in tests.  The
sm_100a variant (``ops.preprocess_nhwc``) fuses /255, the shift, the NCHW→NHWC transpose, the bf16 cast and
the 3→8 channel zero-pad (TMA needs 16-byte rows) into one pass.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch
import torch.nn.functional as F


def crop_pads(height: int, width: int, ratio: float = 0.07) -> Tuple[int, int]:
    return int(height * ratio), int(width * ratio)


def sample_shift(height: int, width: int, ratio: float = 0.07, generator: Optional[torch.Generator] = None):
    ud, lr = crop_pads(height, width, ratio)
    dh = int(torch.randint(0, 2 * ud + 1, size=[], generator=generator))
    dw = int(torch.randint(0, 2 * lr + 1, size=[], generator=generator))
    return dh, dw


def convert_dtype_and_crop_images(images: torch.Tensor, ratio: float = 0.07,
                                  shift: Optional[Tuple[int, int]] = None) -> torch.Tensor:
    """``images`` ``[B, 3, H, W]`` uint8 (0..255) or float (0..1) → float32 ``[B, 3, H, W]``."""
    if images.dtype == torch.uint8:
        images = images / 255.0
    images = images.to(torch.float32)
    _, _, h, w = images.shape
    ud, lr = crop_pads(h, w, ratio)
    dh, dw = shift if shift is not None else sample_shift(h, w, ratio)
    padded = F.pad(images, (lr, lr, ud, ud))
    return padded[..., dh:dh + h, dw:dw + w]
