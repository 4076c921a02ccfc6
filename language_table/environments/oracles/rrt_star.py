"""RRT* in the plane with disc obstacles (reference ``environments/oracles/rrt_star.py:88``: pure-python RRT* used by the
push oracle to route a block, and then the effector, around the other blocks)."""
import math
from typing import List, Optional, Sequence, Tuple

import numpy as np


# ------------------------------------------------------------------------------------------------ geometric predicates
def get_dist(start, end) -> float:
    return float(np.hypot(end[0] - start[0], end[1] - start[1]))


def segment_distances(a, b, centres) -> np.ndarray:
    """Distance of every centre ``[n, 2]`` to the closed segment a–b (closest point, not sampled)."""
    a, b, centres = np.asarray(a, float), np.asarray(b, float), np.asarray(centres, float).reshape(-1, 2)
    d = b - a
    dd = float(d @ d)
    t = np.clip(((centres - a) @ d) / dd, 0.0, 1.0) if dd > 0 else np.zeros(len(centres))
    return np.linalg.norm(a[None] + t[:, None] * d[None] - centres, axis=-1)


def inside_obstacles(node, delta, obstacles) -> bool:
    """Point within ``radius + delta`` of any disc ``(x, y, radius)`` (reference ``rrt_star.py:298``)."""
    obs = np.asarray(obstacles, float).reshape(-1, 3)
    return bool(len(obs)) and bool((np.hypot(obs[:, 0] - node[0], obs[:, 1] - node[1]) <= obs[:, 2] + delta).any())


def inside_boundary(node, delta, obs_boundary) -> bool:
    """Point inside any wall rectangle ``(x, y, w, h)`` grown by ``delta`` on every side (``rrt_star.py:305``)."""
    for x, y, w, h in obs_boundary:
        if 0 <= node[0] - (x - delta) <= w + 2 * delta and 0 <= node[1] - (y - delta) <= h + 2 * delta:
            return True
    return False


def is_inside_obstacles_or_boundary(node, delta, obstacles, obs_boundary) -> bool:
    return inside_obstacles(node, delta, obstacles) or inside_boundary(node, delta, obs_boundary)


def is_intersect_circle(o, d, a, r, delta) -> bool:
    """Does the segment ``o + t d, t in [0, 1]`` come within ``r + delta`` of centre ``a`` at its foot point?  Like the
    reference's test (``rrt_star.py:313``) the end points are the caller's business: a foot point outside the segment
    reports no hit."""
    o, d, a = np.asarray(o, float), np.asarray(d, float), np.asarray(a, float)
    dd = float(d @ d)
    if dd == 0:
        return False
    t = float((a - o) @ d) / dd
    return 0 <= t <= 1 and float(np.linalg.norm(o + t * d - a)) <= r + delta


def get_ray(start, end):
    return [start[0], start[1]], [end[0] - start[0], end[1] - start[1]]


def is_collision(start, end, delta, obstacles, obs_boundary) -> bool:
    """End points inside a grown disc or wall, or the segment passing within ``radius + delta`` of a disc
    (``rrt_star.py:275``)."""
    if is_inside_obstacles_or_boundary(start, delta, obstacles, obs_boundary) or \
            is_inside_obstacles_or_boundary(end, delta, obstacles, obs_boundary):
        return True
    obs = np.asarray(obstacles, float).reshape(-1, 3)
    return bool(len(obs)) and bool((segment_distances(start, end, obs[:, :2]) <= obs[:, 2] + delta).any())


def get_shortest_path_no_collisions(xy_start, xy_goal, x_range, y_range, obstacle_xy, obstacle_widths, delta, step_length,
                                    goal_sample_rate, search_radius, iter_max, boundary_width=0.01, plot=False,
                                    raise_error_on_plan_failure=False, rng=None):
    """The reference's planner entry point (``rrt_star.py:25-85``): discs ``(x, y, width)`` grown by ``delta``, the
    rectangle ``x_range × y_range`` as the free region, and a ``(path, success)`` result whose path runs from the goal
    back to the start; a failed search returns the straight line ``[goal, start]`` (or raises on request) so the caller
    can "just try it" and re-plan later.  ``boundary_width`` is accepted for the signature: the walls here are the
    sampling rectangle itself, not four thin rectangles.  ``plot=True`` writes ``plot.DEBUG_IMAGE_PATH``."""
    obstacles = [(x, y, w + delta) for (x, y), w in zip(obstacle_xy, obstacle_widths)]
    bounds = ((x_range[0], y_range[0]), (x_range[1], y_range[1]))
    planner = RRTStar(xy_start, xy_goal, obstacles, bounds, rng=rng, expand_dis=step_length, goal_sample_rate=goal_sample_rate,
                      max_iter=iter_max, connect_circle_dist=search_radius, search_until_max_iter=True)
    path = planner.plan()
    if plot:
        from language_table.environments.oracles import plot as plot_module
        walls = [(x_range[0], y_range[0], x_range[1] - x_range[0], y_range[1] - y_range[0])]
        plotter = plot_module.PlotRRT(xy_start, xy_goal, walls, obstacles)
        plotter.animation(planner, path, "rrt*")
        plotter.save_debug_image()
    if path is None:
        if raise_error_on_plan_failure:
            raise ValueError("Could not find path! Consider retuning RRT-*.")
        return [np.asarray(xy_goal, float), np.asarray(xy_start, float)], False
    return [p for p in path[::-1]], True


class RRTStar:
    def __init__(self, start, goal, obstacles: Sequence[Tuple[float, float, float]], bounds, rng=None, expand_dis=0.03,
                 path_resolution=0.005, goal_sample_rate=0.1, max_iter=1024, connect_circle_dist=0.2, search_until_max_iter=False):
        self.start, self.goal = np.asarray(start, float), np.asarray(goal, float)
        self.obstacles = np.asarray(obstacles, float).reshape(-1, 3)
        (self.xmin, self.ymin), (self.xmax, self.ymax) = bounds
        self.rng = rng or np.random.RandomState(0)
        self.expand_dis, self.res, self.goal_rate = expand_dis, path_resolution, goal_sample_rate
        self.max_iter, self.connect_dist, self.search_all = max_iter, connect_circle_dist, search_until_max_iter
        self.nodes: List[np.ndarray] = []
        self.parent: List[int] = []
        self.cost: List[float] = []

    def _free(self, a, b) -> bool:
        if len(self.obstacles) == 0:
            return True
        return bool((segment_distances(a, b, self.obstacles[:, :2]) > self.obstacles[:, 2]).all())

    def _sample(self):
        if self.rng.rand() < self.goal_rate:
            return self.goal.copy()
        return np.array([self.rng.uniform(self.xmin, self.xmax), self.rng.uniform(self.ymin, self.ymax)])

    def plan(self) -> Optional[np.ndarray]:
        self.nodes, self.parent, self.cost = [self.start], [-1], [0.0]
        best_goal = None
        for it in range(self.max_iter):
            rnd = self._sample()
            pts = np.asarray(self.nodes)
            near_i = int(np.argmin(np.linalg.norm(pts - rnd, axis=1)))
            direction = rnd - pts[near_i]
            dist = np.linalg.norm(direction)
            if dist < 1e-9:
                continue
            new = pts[near_i] + direction / dist * min(self.expand_dis, dist)
            if not self._free(pts[near_i], new):
                continue
            n = len(self.nodes) + 1
            radius = min(self.connect_dist * math.sqrt(math.log(n) / n), self.expand_dis * 2.5)
            neigh = np.where(np.linalg.norm(pts - new, axis=1) <= max(radius, self.expand_dis))[0]
            parent, cost = near_i, self.cost[near_i] + np.linalg.norm(new - pts[near_i])
            for j in neigh:                                     # choose the cheapest collision-free parent
                c = self.cost[j] + np.linalg.norm(new - pts[j])
                if c < cost and self._free(pts[j], new):
                    parent, cost = int(j), c
            self.nodes.append(new); self.parent.append(parent); self.cost.append(cost)
            me = len(self.nodes) - 1
            for j in neigh:                                     # rewire
                c = cost + np.linalg.norm(new - pts[j])
                if c < self.cost[j] and self._free(new, pts[j]):
                    self.parent[j], self.cost[j] = me, c
            if np.linalg.norm(new - self.goal) <= self.expand_dis and self._free(new, self.goal):
                total = cost + np.linalg.norm(new - self.goal)
                if best_goal is None or total < best_goal[1]:
                    best_goal = (me, total)
                if not self.search_all:
                    break
        if best_goal is None:
            return None
        path, i = [self.goal], best_goal[0]
        while i != -1:
            path.append(self.nodes[i])
            i = self.parent[i]
        return np.asarray(path[::-1])
