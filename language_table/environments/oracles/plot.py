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


DEBUG_IMAGE_PATH = "/tmp/rrt_debug_img.png"


class PlotRRT:
    """The reference's debug plotter (``plot.py:28-114``) on the numpy rasteriser above: same constructor and the same
    ``animation`` → ``save_debug_image`` sequence, a PNG at ``image_path`` as the product.

    ``obs_boundary``: wall rectangles ``(x, y, w, h)`` — their hull is the drawn region; ``obs_circle``: discs
    ``(x, y, r)``.  ``nodelist`` is either a list of tree edges ``((x0, y0), (x1, y1))`` or a planner exposing
    ``nodes`` / ``parent`` (``rrt_star.RRTStar``)."""

    def __init__(self, x_start, x_goal, obs_boundary, obs_circle, image_path=DEBUG_IMAGE_PATH, size: int = 400):
        self._x_start, self._x_goal = x_start, x_goal
        self._obs_bound, self._obs_circle = list(obs_boundary), list(obs_circle)
        self._image_path, self._size = image_path, size
        self.fig = None
        self._edges, self._path, self.name = [], None, ""

    def _bounds(self):
        if not self._obs_bound:
            pts = np.asarray([self._x_start, self._x_goal] + [c[:2] for c in self._obs_circle], float)
            return pts.min(0) - 0.1, pts.max(0) + 0.1
        r = np.asarray(self._obs_bound, float)
        return np.array([r[:, 0].min(), r[:, 1].min()]), np.array([(r[:, 0] + r[:, 2]).max(), (r[:, 1] + r[:, 3]).max()])

    def plot_grid(self, name):
        self.name = name
        self._render()

    def plot_visited(self, nodelist, animation=False):
        if hasattr(nodelist, "nodes") and hasattr(nodelist, "parent"):
            self._edges = [(nodelist.nodes[p], n) for n, p in zip(nodelist.nodes, nodelist.parent) if p >= 0]
        else:
            self._edges = list(nodelist or [])
        self._render()

    def plot_path(self, path):
        self._path = None if path is None or len(path) == 0 else [np.asarray(p, float) for p in path]
        self._render()

    def animation(self, nodelist, path, name, animation=False):
        self.name = name
        self.plot_visited(nodelist, animation)
        self.plot_path(path)

    def _render(self):
        lo, hi = self._bounds()
        self.fig = plot_plan(self._path, self._obs_circle, self._x_start, self._x_goal, bounds=(lo, hi), edges=self._edges,
                             size=self._size)
        return self.fig

    def fig_to_array(self, figure=None):
        return self._render() if figure is None else np.asarray(figure)

    def save_debug_image(self):
        from rt1b200.utils.video import write_apng
        return write_apng(self._image_path, self.fig_to_array()[None])
