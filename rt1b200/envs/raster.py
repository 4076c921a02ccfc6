"""Frame rendering for the synthetic board: a CUDA z-buffer rasteriser and its CPU restatement.

``render`` dispatches to the CUDA kernel (``csrc/raster.cu``) for CUDA inputs and to :func:`render_reference` — the
same screen-space edge-function / 1-over-z test written with numpy — otherwise.  The CPU version is the oracle for
the kernel test and the timed CPU baseline (the reference's TinyRenderer and any graphics-API renderer are unavailable
on this box).  ``crop_resize`` is the fused centre-crop + bilinear resize that produces the policy's image input
(reference ``language_table/eval/wrappers.py:88-96``).
"""
from __future__ import annotations

import numpy as np
import torch

from language_table.environments import constants as C

from ..ops import _lib
from ..ops._lib import call, f32, i32, ptr, stream
from .scene import BACKGROUND_RGB


def render_reference(tris: np.ndarray, ntris: np.ndarray, H: int = C.IMAGE_HEIGHT, W: int = C.IMAGE_WIDTH) -> np.ndarray:
    """tris [E,T,12], ntris [E] → uint8 [E,H,W,3] (nearest triangle per pixel, sampled at pixel centres)."""
    tris = np.asarray(tris, np.float32)
    E = tris.shape[0]
    out = np.empty((E, H, W, 3), np.uint8)
    ys, xs = np.mgrid[0:H, 0:W]
    cx, cy = xs + 0.5, ys + 0.5
    for e in range(E):
        best = np.zeros((H, W), np.float32)
        img = np.empty((H, W, 3), np.float32)
        img[:] = BACKGROUND_RGB
        for t in tris[e, :int(ntris[e])]:
            x0, y0, w0, x1, y1, w1, x2, y2, w2 = t[:9]
            area = (x1 - x0) * (y2 - y0) - (x2 - x0) * (y1 - y0)
            if abs(area) < 1e-9:
                continue
            lo_x, hi_x = int(max(0, np.floor(min(x0, x1, x2)) - 1)), int(min(W, np.ceil(max(x0, x1, x2)) + 1))
            lo_y, hi_y = int(max(0, np.floor(min(y0, y1, y2)) - 1)), int(min(H, np.ceil(max(y0, y1, y2)) + 1))
            if lo_x >= hi_x or lo_y >= hi_y:
                continue
            px, py = cx[lo_y:hi_y, lo_x:hi_x], cy[lo_y:hi_y, lo_x:hi_x]
            b0 = ((x1 - px) * (y2 - py) - (x2 - px) * (y1 - py)) / area
            b1 = ((x2 - px) * (y0 - py) - (x0 - px) * (y2 - py)) / area
            b2 = ((x0 - px) * (y1 - py) - (x1 - px) * (y0 - py)) / area
            w = b0 * w0 + b1 * w1 + b2 * w2
            hit = (b0 >= 0) & (b1 >= 0) & (b2 >= 0) & (w > best[lo_y:hi_y, lo_x:hi_x])
            best[lo_y:hi_y, lo_x:hi_x][hit] = w[hit]
            img[lo_y:hi_y, lo_x:hi_x][hit] = t[9:12]
        out[e] = np.clip(img, 0, 255).astype(np.uint8)
    return out


def render(tris: torch.Tensor, ntris: torch.Tensor, H: int = C.IMAGE_HEIGHT, W: int = C.IMAGE_WIDTH) -> torch.Tensor:
    if tris.is_cuda and _lib.available():
        E, T = tris.shape[0], tris.shape[1]
        out = torch.empty(E, H, W, 3, dtype=torch.uint8, device=tris.device)
        call("rt1_raster", ptr(tris.contiguous()), ptr(ntris.to(torch.int32).contiguous()), i32(T), ptr(out), i32(E), i32(H), i32(W),
             f32(BACKGROUND_RGB[0]), f32(BACKGROUND_RGB[1]), f32(BACKGROUND_RGB[2]), stream())
        return out
    return torch.from_numpy(render_reference(tris.cpu().numpy(), ntris.cpu().numpy(), H, W)).to(tris.device)


def crop_resize_reference(frames: torch.Tensor, out_h: int, out_w: int, crop: float = 0.95) -> torch.Tensor:
    """uint8 [E,H,W,3] → float32 [E,3,out_h,out_w] in [0,1]; same sampling rule as the kernel (half-pixel centres)."""
    E, H, W, _ = frames.shape
    import math
    ch, cw = float(math.floor(H * crop)), float(math.floor(W * crop))           # whole-pixel window, like the reference's
    y0, x0 = float(math.floor((H - H * crop) / 2)), float(math.floor((W - W * crop) / 2))   # crop_test_image
    fy = y0 + (torch.arange(out_h, dtype=torch.float32) + 0.5) * (ch / out_h) - 0.5
    fx = x0 + (torch.arange(out_w, dtype=torch.float32) + 0.5) * (cw / out_w) - 0.5
    ylo, xlo = int(y0), int(x0)
    yhi, xhi = min(H - 1, ylo + int(ch) - 1), min(W - 1, xlo + int(cw) - 1)      # samples stay inside the crop window
    iy0 = fy.floor().clamp(ylo, yhi).long(); ix0 = fx.floor().clamp(xlo, xhi).long()
    iy1 = (iy0 + 1).clamp(max=yhi); ix1 = (ix0 + 1).clamp(max=xhi)
    ty = (fy - iy0).clamp(0, 1)[None, :, None, None]; tx = (fx - ix0).clamp(0, 1)[None, None, :, None]
    f = frames.float().cpu()
    v = (f[:, iy0][:, :, ix0] * (1 - tx) + f[:, iy0][:, :, ix1] * tx) * (1 - ty) + \
        (f[:, iy1][:, :, ix0] * (1 - tx) + f[:, iy1][:, :, ix1] * tx) * ty
    return (v / 255.0).permute(0, 3, 1, 2).contiguous().to(frames.device)


def crop_resize(frames: torch.Tensor, out_h: int, out_w: int, crop: float = 0.95) -> torch.Tensor:
    if frames.is_cuda and _lib.available():
        E, H, W, _ = frames.shape
        out = torch.empty(E, 3, out_h, out_w, dtype=torch.float32, device=frames.device)
        call("rt1_crop_resize", ptr(frames.contiguous()), ptr(out), i32(E), i32(H), i32(W), i32(out_h), i32(out_w), f32(crop), stream())
        return out
    return crop_resize_reference(frames, out_h, out_w, crop)
