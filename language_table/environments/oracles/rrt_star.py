"""RRT* in the plane with disc obstacles (reference ``environments/oracles/rrt_star.py:88``: pure-python RRT* used by the
push oracle to route a block, and then the effector, around the other blocks)."""
import math
from typing import List, Optional, Sequence, Tuple

import numpy as np


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
        n = max(2, int(np.linalg.norm(b - a) / self.res) + 1)
        pts = a[None] + (b - a)[None] * np.linspace(0, 1, n)[:, None]
        d = np.linalg.norm(pts[:, None, :] - self.obstacles[None, :, :2], axis=-1)
        return bool((d > self.obstacles[None, :, 2]).all())

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
