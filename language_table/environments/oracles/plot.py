"""Debug picture of an RRT* plan (reference ``language_table/environments/oracles/plot.py:28-114`` draws the tree, the
obstacles and the path with matplotlib, which is not installed offline): a small numpy rasteriser writing a PPM file."""
from typing import Iterable, Optional, Sequence, Tuple

import numpy as np


def _to_px(p, lo, hi, size):
    return np.clip(((np.asarray(p, dtype=np.float64) - lo) / (hi - lo) * (size - 1)).astype(int), 0, size - 1)


def _line(img, a, b, color):
    n = int(max(abs(b[0] - a[0]), abs(b[1] - a[1]))) + 1
    xs, ys = np.linspace(a[0], b[0], n).astype(int), np.linspace(a[1], b[1], n).astype(int)
    img[ys, xs] = color


def _disc(img, c, r, color):
    yy, xx = np.mgrid[0:img.shape[0], 0:img.shape[1]]
    img[(xx - c[0]) ** 2 + (yy - c[1]) ** 2 <= r * r] = color


def plot_plan(path: Optional[Sequence[Sequence[float]]], obstacles: Iterable[Tuple[float, float, float]], start, goal,
              bounds=((0.15, -0.3048), (0.6, 0.3048)), edges: Optional[Iterable] = None, size: int = 400,
              filename: Optional[str] = None) -> np.ndarray:
    """``obstacles``: (x, y, radius); ``edges``: tree edges ((x0, y0), (x1, y1)); returns the uint8 [size, size, 3] image."""
    lo, hi = np.asarray(bounds[0], dtype=np.float64), np.asarray(bounds[1], dtype=np.float64)
    img = np.full((size, size, 3), 255, np.uint8)
    scale = (size - 1) / float(np.max(hi - lo))
    for ox, oy, r in obstacles:
        _disc(img, _to_px((ox, oy), lo, hi, size), max(1, int(r * scale)), (120, 120, 120))
    for a, b in (edges or []):
        _line(img, _to_px(a, lo, hi, size), _to_px(b, lo, hi, size), (170, 200, 255))
    if path is not None and len(path) > 1:
        for a, b in zip(path[:-1], path[1:]):
            _line(img, _to_px(a, lo, hi, size), _to_px(b, lo, hi, size), (220, 30, 30))
    _disc(img, _to_px(start, lo, hi, size), 4, (30, 160, 30))
    _disc(img, _to_px(goal, lo, hi, size), 4, (30, 30, 200))
    if filename:
        with open(filename, "wb") as f:
            f.write(f"P6 {size} {size} 255\n".encode() + img.tobytes())
    return img
