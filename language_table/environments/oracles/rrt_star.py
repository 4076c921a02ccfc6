"""RRT* in the plane with disc obstacles (reference ``environments/oracles/rrt_star.py:88``: pure-python RRT* used by the
push oracle to route a block, and then the effector, around the other blocks)."""
import collections
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


class Node:
    """A point with a parent link — the record the reference's planner is built on (``rrt_star.py:335-341``).  The
    planner below keeps its tree in flat lists instead; ``RRTStar.node(i)`` builds these on request."""

    def __init__(self, n):
        self.x, self.y = float(n[0]), float(n[1])
        self.parent = None


class FIFOQueue:
    """First in, first out (``rrt_star.py:344-357``); used to walk a re-parented subtree breadth first."""

    def __init__(self):
        self.queue = collections.deque()

    def empty(self) -> bool:
        return not self.queue

    def put(self, node) -> None:
        self.queue.append(node)

    def get(self):
        return self.queue.popleft()


class RRTStar:
    """RRT* over a flat tree: ``nodes[i]`` (a point), ``parent[i]`` (an index, -1 at the root), ``cost[i]`` (path length
    from the root), ``children[i]``.  One iteration = sample (``get_random_node``) → nearest tree node
    (``nearest_neighbor``) → step towards the sample (``new_state``) → neighbours inside the shrinking search ball that can
    be reached without collision (``find_nearest_neighbors``) → cheapest parent among them (``choose_parent``) → re-parent
    neighbours that get cheaper through the new node and push the saving down their subtrees (``rewire``,
    ``update_cost``).  ``search_goal_parent`` picks the cheapest node one step from the goal, ``extract_path`` walks back
    to the root.  The step names are the reference's (``rrt_star.py:122-272``); obstacles are discs ``(x, y, radius)``
    with any clearance already added, the free region is the rectangle ``bounds``."""

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
        self.cost_to: List[float] = []
        self.children: List[List[int]] = []
        self.path: Optional[np.ndarray] = None
        self.success = False

    # ------------------------------------------------------------------------------------------------ geometry
    def _free(self, a, b) -> bool:
        if len(self.obstacles) == 0:
            return True
        return bool((segment_distances(a, b, self.obstacles[:, :2]) > self.obstacles[:, 2]).all())

    @staticmethod
    def get_distance_and_angle(node_start, node_end):
        dx, dy = node_end[0] - node_start[0], node_end[1] - node_start[1]
        return math.hypot(dx, dy), math.atan2(dy, dx)

    @staticmethod
    def nearest_neighbor(node_list, n) -> int:
        """Index of the point of ``node_list`` closest to ``n``."""
        return int(np.argmin(np.linalg.norm(np.asarray(node_list, float) - np.asarray(n, float), axis=1)))

    def node(self, i: int) -> Node:
        n = Node(self.nodes[i])
        n.parent = None if self.parent[i] < 0 else Node(self.nodes[self.parent[i]])
        return n

    # ------------------------------------------------------------------------------------------------ one iteration
    def get_random_node(self, goal_sample_rate: Optional[float] = None) -> np.ndarray:
        rate = self.goal_rate if goal_sample_rate is None else goal_sample_rate
        if self.rng.rand() < rate:
            return self.goal.copy()
        return np.array([self.rng.uniform(self.xmin, self.xmax), self.rng.uniform(self.ymin, self.ymax)])

    def new_state(self, node_start, node_goal) -> Optional[np.ndarray]:
        """The point at most one step from ``node_start`` towards ``node_goal`` (``None`` when they coincide)."""
        dist, theta = self.get_distance_and_angle(node_start, node_goal)
        if dist < 1e-9:
            return None
        step = min(self.expand_dis, dist)
        return np.array([node_start[0] + step * math.cos(theta), node_start[1] + step * math.sin(theta)])

    def find_nearest_neighbors(self, node_new) -> np.ndarray:
        """Tree nodes inside the search ball — radius ``connect_circle_dist · sqrt(log n / n)``, capped at 2.5 steps and at
        least one step — from which ``node_new`` is reachable in a straight line."""
        n = len(self.nodes) + 1
        radius = max(min(self.connect_dist * math.sqrt(math.log(n) / n), self.expand_dis * 2.5), self.expand_dis)
        pts = np.asarray(self.nodes)
        near = np.where(np.linalg.norm(pts - node_new, axis=1) <= radius)[0]
        return np.array([j for j in near if self._free(pts[j], node_new)], dtype=np.int64)

    def cost(self, node_p: int) -> float:
        """Path length from the root to node ``node_p``."""
        return self.cost_to[node_p]

    def get_new_cost(self, node_start: int, node_end) -> float:
        """Cost of ``node_end`` (a point) if reached through tree node ``node_start``."""
        return self.cost_to[node_start] + float(np.linalg.norm(np.asarray(node_end) - self.nodes[node_start]))

    def choose_parent(self, node_new, neighbor_index) -> Tuple[int, float]:
        costs = [self.get_new_cost(int(j), node_new) for j in neighbor_index]
        k = int(np.argmin(costs))
        return int(neighbor_index[k]), costs[k]

    def _attach(self, point, parent: int, cost: float) -> int:
        self.nodes.append(np.asarray(point, float))
        self.parent.append(parent)
        self.cost_to.append(cost)
        self.children.append([])
        if parent >= 0:
            self.children[parent].append(len(self.nodes) - 1)
        return len(self.nodes) - 1

    def update_cost(self, parent_node: int) -> None:
        """Breadth-first refresh of the costs below ``parent_node`` after it was given a cheaper parent."""
        queue = FIFOQueue()
        queue.put(parent_node)
        while not queue.empty():
            i = queue.get()
            for c in self.children[i]:
                self.cost_to[c] = self.get_new_cost(i, self.nodes[c])
                queue.put(c)

    def rewire(self, node_new: int, neighbor_index) -> None:
        for j in neighbor_index:
            j = int(j)
            c = self.get_new_cost(node_new, self.nodes[j])
            if c < self.cost_to[j] and j != self.parent[node_new]:
                self.children[self.parent[j]].remove(j)
                self.parent[j], self.cost_to[j] = node_new, c
                self.children[node_new].append(j)
                self.update_cost(j)

    def search_goal_parent(self) -> Optional[int]:
        """The tree node within one step of the goal, with a free straight line to it, that gives the shortest total path."""
        pts = np.asarray(self.nodes)
        close = np.where(np.linalg.norm(pts - self.goal, axis=1) <= self.expand_dis)[0]
        best, best_cost = None, float("inf")
        for j in close:
            c = self.get_new_cost(int(j), self.goal)
            if c < best_cost and self._free(pts[j], self.goal):
                best, best_cost = int(j), c
        return best

    def extract_path(self, node_end: int) -> np.ndarray:
        """Start → ... → ``node_end`` → goal."""
        path, i = [self.goal], node_end
        while i != -1:
            path.append(self.nodes[i])
            i = self.parent[i]
        return np.asarray(path[::-1])

    # ------------------------------------------------------------------------------------------------ the search
    def plan(self) -> Optional[np.ndarray]:
        self.nodes, self.parent, self.cost_to, self.children = [], [], [], []
        self._attach(self.start, -1, 0.0)
        self.path, self.success = None, False
        for _ in range(self.max_iter):
            rnd = self.get_random_node()
            near_i = self.nearest_neighbor(self.nodes, rnd)
            new = self.new_state(self.nodes[near_i], rnd)
            if new is None or not self._free(self.nodes[near_i], new):
                continue
            neighbors = self.find_nearest_neighbors(new)
            if len(neighbors) == 0:
                continue
            parent, cost = self.choose_parent(new, neighbors)
            me = self._attach(new, parent, cost)
            self.rewire(me, neighbors)
            if not self.search_all and np.linalg.norm(new - self.goal) <= self.expand_dis and self._free(new, self.goal):
                break
        end = self.search_goal_parent()
        if end is None:
            return None
        self.path, self.success = self.extract_path(end), True
        return self.path
