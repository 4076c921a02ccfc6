"""Device-side half of the training input pipeline.

The reference's loader workers crop 95 % of every frame at a random offset, resize it to 456 x 256 with PIL and convert it
to float32 — per frame, on the CPU (``load_np_dataset.py:8-39``); at B200 training rates (hundreds of samples/s per GPU, six
frames each) that is thousands of PIL resizes per second and an H2D stream of 5.6 MB per sample.  Here a worker only SAMPLES
the crop box and ships the raw uint8 frame (:class:`RawFramesWithBoxes`, 0.17 MB per 180 x 320 frame); the crop, the bilinear
resize and (inside the model) the /255 + pad-shift + NHWC conversion run as kernels on the GPU (:class:`GpuCropResize`).
"""
from __future__ import annotations

from typing import Dict, Optional, Tuple

import numpy as np
import torch

from ..ops import _lib
from ..ops._lib import call, i32, ptr, stream


def sample_crop_box(raw_h: int, raw_w: int, factor: Optional[float], rng: np.random.Generator) -> np.ndarray:
    """(top, left, height, width) of the reference's random crop: a window of ``factor`` of each side at a uniformly random
    integer offset (``load_np_dataset.py:17-31``); the whole frame when ``factor`` is None."""
    if factor is None:
        return np.array([0.0, 0.0, raw_h, raw_w], np.float32)
    sh, sw = raw_h * factor, raw_w * factor
    top = int(rng.integers(0, int(raw_h - sh + 1)))
    left = int(rng.integers(0, int(raw_w - sw + 1)))
    return np.array([top, left, sh, sw], np.float32)


class RawFramesWithBoxes:
    """Dataset-side transform for a WINDOW of frames: keeps them raw (uint8 HWC) and attaches one crop box per frame."""
    windowed = True

    def __init__(self, random_crop_factor: Optional[float] = None, rng: Optional[np.random.Generator] = None):
        self.random_crop_factor = random_crop_factor
        self.rng = rng or np.random.default_rng()

    def __call__(self, frames: np.ndarray) -> Dict[str, torch.Tensor]:
        t, h, w, _ = frames.shape
        boxes = np.stack([sample_crop_box(h, w, self.random_crop_factor, self.rng) for _ in range(t)])
        return {"image_raw": torch.from_numpy(np.ascontiguousarray(frames)), "crop_boxes": torch.from_numpy(boxes)}


def crop_resize_boxes_reference(frames: torch.Tensor, boxes: torch.Tensor, out_hw: Tuple[int, int], as_uint8: bool = True) -> torch.Tensor:
    """PyTorch statement of ``rt1_crop_resize_boxes`` (the oracle of its test and the CPU path)."""
    E, H, W, _ = frames.shape
    OH, OW = out_hw
    f = frames.float()
    out = torch.empty(E, 3, OH, OW, dtype=torch.float32, device=frames.device)
    oy = torch.arange(OH, dtype=torch.float32, device=frames.device) + 0.5
    ox = torch.arange(OW, dtype=torch.float32, device=frames.device) + 0.5
    for e in range(E):
        top, left, bh, bw = (float(v) for v in boxes[e])
        fy, fx = top + oy * (bh / OH) - 0.5, left + ox * (bw / OW) - 0.5
        ylo, yhi = max(0, int(np.floor(top))), min(H - 1, int(np.ceil(top + bh)) - 1)
        xlo, xhi = max(0, int(np.floor(left))), min(W - 1, int(np.ceil(left + bw)) - 1)
        iy0 = fy.floor().clamp(ylo, yhi).long(); ix0 = fx.floor().clamp(xlo, xhi).long()
        iy1 = (iy0 + 1).clamp(max=yhi); ix1 = (ix0 + 1).clamp(max=xhi)
        ty = (fy - iy0).clamp(0, 1)[:, None, None]; tx = (fx - ix0).clamp(0, 1)[None, :, None]
        img = f[e]
        v = (img[iy0][:, ix0] * (1 - tx) + img[iy0][:, ix1] * tx) * (1 - ty) + (img[iy1][:, ix0] * (1 - tx) + img[iy1][:, ix1] * tx) * ty
        out[e] = v.permute(2, 0, 1)
    return (out + 0.5).floor().clamp(0, 255).to(torch.uint8) if as_uint8 else out / 255.0


def crop_resize_boxes(frames: torch.Tensor, boxes: torch.Tensor, out_hw: Tuple[int, int], as_uint8: bool = True) -> torch.Tensor:
    """``frames`` uint8 ``[E, H, W, 3]``, ``boxes`` f32 ``[E, 4]`` → ``[E, 3, OH, OW]`` (uint8, or f32 in [0, 1])."""
    if frames.is_cuda and _lib.available():
        E, H, W, _ = frames.shape
        frames = frames.contiguous()
        boxes = boxes.to(frames.device, torch.float32).contiguous()
        out = torch.empty(E, 3, out_hw[0], out_hw[1], dtype=torch.uint8 if as_uint8 else torch.float32, device=frames.device)
        call("rt1_crop_resize_boxes", ptr(frames), ptr(boxes), ptr(out), i32(as_uint8), i32(E), i32(H), i32(W), i32(out_hw[0]),
             i32(out_hw[1]), stream())
        return out
    return crop_resize_boxes_reference(frames, boxes, out_hw, as_uint8)


class GpuCropResize:
    """Batch transform applied by the trainer once a batch is on the device: ``image_raw`` + ``crop_boxes`` → ``image``."""

    def __init__(self, height: int, width: int):
        self.out_hw = (height, width)

    def __call__(self, batch: Dict) -> Dict:
        obs = batch.get("train_observation", {})
        if "image_raw" not in obs:
            return batch
        raw, boxes = obs["image_raw"], obs["crop_boxes"]
        b, t = raw.shape[:2]
        img = crop_resize_boxes(raw.reshape(b * t, *raw.shape[2:]), boxes.reshape(b * t, 4), self.out_hw, as_uint8=True)
        new_obs = {k: v for k, v in obs.items() if k not in ("image_raw", "crop_boxes")}
        new_obs["image"] = img.view(b, t, 3, *self.out_hw)
        return {**batch, "train_observation": new_obs}
