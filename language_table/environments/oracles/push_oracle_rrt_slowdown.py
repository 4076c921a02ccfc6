"""Obstacle-aware pushing oracle (reference ``environments/oracles/push_oracle_rrt_slowdown.py:95``).

Two planning levels, both RRT* in the plane (``rrt_star.RRTStar``):

* **block plan** (``get_plan``, reference ``:177-282``) — the route of the pushed block to its goal around the other blocks
  (small collision discs, the planner only has to keep block centres apart); when no route to the goal exists and the goal sits
  next to another block (block-to-block-relative tasks), routes to points backed off along the line away from that block are
  tried and the true goal is appended.  ``get_plan`` is also what the evaluator uses to reject unsolvable initial boards
  (``eval/main_rt1.py:162-172``);
* **end-effector plan** (``use_ee_planner``, reference ``:467-535``) — the free-space route of the effector from where it is to
  the pre-push pose behind the block, around ALL blocks (blocks already touching the pre-push point other than the pushed one
  are ignored), with pre-push offsets of 5 cm and, failing that, 6/7/8 cm.  The subgoals are followed open-loop; the last leg
  tracks the closed-loop pre-push point because a light touch moves the block.
* **potential-field fallback** (reference ``:548-600``) — when no effector route is found (typically right after a push, with
  the effector still against a block): repulsion from blocks closer than 5.9 cm, attraction to the pre-push point; planning
  is retried on later steps.

Pushing itself is the oriented-push state machine of the base class, aimed at the current block subgoal; steps are shortened
near the goal by the reference's slowdown table (``:322-331``).
"""
import collections
from typing import List, Optional

import numpy as np

from language_table.environments import constants
from language_table.environments.oracles import oriented_push_oracle, rrt_star

# planning constants of the reference (``:28-75``)
BLOCK_DIAMETER = 0.015                      # collision disc put around each obstacle block for the block plan
RRT_COLLISION_THRESHOLD = 0.015             # extra clearance of the pushed block
ADVANCE_TO_NEXT_SUBGOAL_THRESHOLD = 0.025
PREBLOCK_OFFSET = 0.05
EE_BACKOFF_OFFSETS = (0.06, 0.07, 0.08)
RRT_STEP_LENGTH = 0.05
RRT_GOAL_SAMPLE_RATE = 0.1
RRT_SEARCH_RADIUS = 0.5                     # neighbourhood scale of the choose-parent / rewire steps
RRT_MAX_ITERS = 1024
REPLAN_IF_FAILURE = True                    # a failed search is retried every RETRY_FOR_NEW_PLAN_EVERY steps
EPS = 1e-5
RETRY_FOR_NEW_PLAN_EVERY = 10
ADVANCE_TO_NEXT_EE_SUBGOAL_THRESHOLD = 0.01
EE_RRT_STEP_LENGTH = 0.025
EE_RRT_DELTA = 0.01
EE_RRT_OBSTACLE_RADIUS = 0.02
EE_RRT_ITER_MAX = 2048
RETRY_FOR_NEW_EE_PLAN_EVERY = 1
EXTRA_BOUNDARY_BUFFER = 0.04
BEYOND_TABLE_THRESHOLD = 2.0
# planning may leave the official workspace on the three sides that are not next to the robot
X_RANGE_RRT = (constants.X_MIN, constants.X_MAX + EXTRA_BOUNDARY_BUFFER)
Y_RANGE_RRT = (constants.Y_MIN - EXTRA_BOUNDARY_BUFFER, constants.Y_MAX + EXTRA_BOUNDARY_BUFFER)
PLAN_BOUNDS = ((X_RANGE_RRT[0], Y_RANGE_RRT[0]), (X_RANGE_RRT[1], Y_RANGE_RRT[1]))
SLOWDOWN_TABLE = ((0.02, 0.2), (0.04, 0.3), (0.06, 0.4), (0.08, 0.5), (0.1, 0.6))      # (distance below, step factor)


def filter_subgoals(path, min_distance: float) -> List[np.ndarray]:
    """Thin a start→goal list of points so that consecutive kept points are at least ``min_distance`` apart; the start
    is always kept and so is the goal."""
    pts = [np.asarray(p, float) for p in path]
    if not pts:
        return []
    keep = [pts[0]]
    for p in pts[1:-1]:
        if np.linalg.norm(p - keep[-1]) >= min_distance:
            keep.append(p)
    if len(pts) > 1:
        keep.append(pts[-1])
    return keep


def slowdown_factor(distance: float) -> float:
    for below, factor in SLOWDOWN_TABLE:
        if distance < below:
            return factor
    return 1.0


def repulsive_push(xy_ee: np.ndarray, obstacle_xy: np.ndarray, r: float = 0.029, s: float = 0.03, beta: float = 500.0) -> np.ndarray:
    """Gradient of the repulsive potential of one block on the effector: zero beyond ``r + s``, linear inside the shell,
    overwhelming once the effector is inside the block's disc."""
    d = xy_ee - obstacle_xy
    dist = float(np.linalg.norm(d))
    if dist > r + s:
        return np.zeros(2)
    direction = d / max(dist, 1e-9)
    if dist < r:
        return direction * 1e9
    return direction * beta * (s + r - dist)


def attractive_pull(xy_ee: np.ndarray, goal_xy: np.ndarray, r: float = 0.03, s: float = 0.03) -> np.ndarray:
    """Attraction towards the pre-push point; the gain grows as the effector gets close (reference ``:577-597``)."""
    d = goal_xy - xy_ee
    dist = float(np.linalg.norm(d))
    direction = d / max(dist, 1e-9)
    if dist > 2 * r:
        return direction * 300.0 * s
    if dist >= r:
        return direction * 550.0 * s
    return direction * 1000.0 * r


class ObstacleOrientedPushOracleBoard2dRRT(oriented_push_oracle.OrientedPushOracle):
    def __init__(self, env, use_ee_planner=True, action_noise_std=0.0, max_step=0.03, slowdown_freespace=False,
                 backoff_subgoal_rrt=True, replan_ee_rrt=True, backoff_ee_rrt=True, filter_ee_obstacle_poses=True, rng=None):
        super().__init__(env, action_noise_std=action_noise_std, max_step=max_step, rng=rng)
        self._use_ee_planner = use_ee_planner
        self._slowdown_freespace = slowdown_freespace
        self._backoff_subgoal_rrt = backoff_subgoal_rrt
        self._replan_ee_rrt = replan_ee_rrt
        self._backoff_ee_rrt = backoff_ee_rrt
        self._filter_ee_obstacle_poses = filter_ee_obstacle_poses
        self.reset()

    def reset(self):
        super().reset()
        self._plan: Optional[collections.deque] = None      # remaining block subgoals, next one first
        self._need_replan = False
        self._replan_counter = 0
        self._ee_plan: Optional[collections.deque] = None   # remaining effector subgoals of the free-space phase
        self._ee_plan_success = False
        self._ee_replan_counter = 0
        self._prev_instruction = None
        self.used_potential_fallback = False

    @property
    def plan_success(self) -> bool:
        return self._plan is not None and not self._need_replan

    # ------------------------------------------------------------------------------------------ scene helpers
    def _task_blocks(self):
        info = getattr(self._env, "_task_info", None)
        block = getattr(info, "block1", None) or getattr(info, "block", None) or getattr(info, "block_target", None)
        return info, block

    def _block_positions(self, state, exclude=()):
        out = []
        for k, v in state.items():
            if k.startswith("block_") and k.endswith("_translation"):
                name = k[len("block_"):-len("_translation")]
                if name in exclude or float(state.get("block_%s_mask" % name, [1.0])[0]) == 0.0:
                    continue
                if float(np.max(v)) >= BEYOND_TABLE_THRESHOLD:       # parked off the table
                    continue
                out.append(np.asarray(v[:2], float))
        return out

    def _rrt(self, start, goal, obstacles, radius, step, max_iter):
        obs = [(float(o[0]), float(o[1]), radius) for o in obstacles]
        planner = rrt_star.RRTStar(start, goal, obs, PLAN_BOUNDS, rng=self._rng, expand_dis=step,
                                   goal_sample_rate=RRT_GOAL_SAMPLE_RATE, max_iter=max_iter)
        return planner.plan()

    # ---------------------------------------------------------------------------------------------- block plan
    def get_plan(self, state) -> bool:
        """Plan the pushed block's route; True when a collision-free route to the goal (or a backed-off goal) exists."""
        info, block = self._task_blocks()
        if block is None:
            return False
        exclude = {block}
        if hasattr(info, "block2"):
            exclude.add(info.block2)                         # block-to-block: the goal block is not an obstacle
        start, goal = self._goal_and_block(state)
        obstacles = [o for o in self._block_positions(state, exclude) if np.linalg.norm(o - goal) > 1e-5]
        radius = BLOCK_DIAMETER + RRT_COLLISION_THRESHOLD
        path = self._rrt(start, goal, obstacles, radius, RRT_STEP_LENGTH, RRT_MAX_ITERS)
        if path is None and self._backoff_subgoal_rrt:
            # the goal may sit right next to another block (relative-location tasks): push to a point a little further out
            near = [o for o in obstacles if np.linalg.norm(o - goal) < 0.1]
            if near:
                away = goal - near[0]
                for scale in (1.1, 1.2, 1.3, 1.4, 1.5):
                    cand = near[0] + away * scale
                    path = self._rrt(start, cand, obstacles, radius, RRT_STEP_LENGTH, RRT_MAX_ITERS)
                    if path is not None:
                        path = np.concatenate([path, goal[None]], 0)
                        break
        success = path is not None
        if not success:
            path = np.stack([start, goal])                  # push straight at it and retry planning later
        self._need_replan = not success
        subgoals = filter_subgoals(list(path), ADVANCE_TO_NEXT_SUBGOAL_THRESHOLD)[1:]        # drop the start
        self._plan = collections.deque(subgoals if subgoals else [goal])
        return success

    # ------------------------------------------------------------------------------------- end-effector plan
    def _get_ee_plan(self, state, block_xy, direction):
        """Route of the effector to the pre-push point around every block; candidates further behind the block are tried
        when the nominal pre-push point starts in collision."""
        xy_ee = np.asarray(state["effector_target_translation"], float)
        pre_block = block_xy - direction * PREBLOCK_OFFSET
        offsets = (PREBLOCK_OFFSET,) + (EE_BACKOFF_OFFSETS if self._backoff_ee_rrt else ())
        path = None
        for off in offsets:
            target = block_xy - direction * off
            obstacles = self._block_positions(state)
            if self._filter_ee_obstacle_poses:
                obstacles = [o for o in obstacles
                             if not (np.linalg.norm(o - target) < 0.05 and np.linalg.norm(o - block_xy) > 1e-6)]
            path = self._rrt(xy_ee, target, obstacles, EE_RRT_OBSTACLE_RADIUS + EE_RRT_DELTA, EE_RRT_STEP_LENGTH, EE_RRT_ITER_MAX)
            if path is not None:
                break
        self._ee_plan_success = path is not None
        pts = filter_subgoals(list(path), ADVANCE_TO_NEXT_EE_SUBGOAL_THRESHOLD)[1:] if path is not None else []
        if not pts or np.linalg.norm(pts[-1] - pre_block) > 1e-5:
            pts.append(pre_block)                           # the route ended at a backed-off point: finish at the real one
        self._ee_plan = collections.deque(pts)

    def _move_to_preblock_potential(self, state, pre_block, block_xy):
        xy_ee = np.asarray(state["effector_target_translation"], float)
        to_pre = pre_block - xy_ee
        if np.linalg.norm(to_pre) < 0.015:
            return to_pre
        delta = attractive_pull(xy_ee, pre_block)
        for o in self._block_positions(state):
            delta = delta + repulsive_push(xy_ee, o)
        return delta

    def _free_space_action(self, state, block_xy, waypoint):
        """Effector motion of the "move to the pre-push pose" phase; returns None once the pose is reached."""
        xy_ee = np.asarray(state["effector_target_translation"], float)
        to_wp = waypoint - block_xy
        direction = to_wp / max(np.linalg.norm(to_wp), 1e-6)
        pre_block = block_xy - direction * PREBLOCK_OFFSET
        if np.linalg.norm(pre_block - xy_ee) < 0.01:
            self._ee_plan = None
            return None
        self._ee_replan_counter += 1
        if self._ee_plan is None or (self._replan_ee_rrt and not self._ee_plan_success
                                     and self._ee_replan_counter % RETRY_FOR_NEW_EE_PLAN_EVERY == 0):
            self._get_ee_plan(state, block_xy, direction)
        if not self._ee_plan_success:
            self.used_potential_fallback = True
            return self._move_to_preblock_potential(state, pre_block, block_xy)
        while len(self._ee_plan) > 1 and np.linalg.norm(self._ee_plan[0] - xy_ee) < ADVANCE_TO_NEXT_EE_SUBGOAL_THRESHOLD:
            self._ee_plan.popleft()
        target = self._ee_plan[0] if len(self._ee_plan) > 1 else pre_block      # last leg: track the live pre-push point
        return target - xy_ee

    # -------------------------------------------------------------------------------------------------- action
    def action(self, state=None):
        state = state or self._env.compute_state()
        instr = state.get("instruction")
        if instr is not None:
            if self._prev_instruction is not None and np.any(np.asarray(instr) != self._prev_instruction):
                self.reset()                                  # a new instruction invalidates both plans
            self._prev_instruction = np.asarray(instr).copy()
        if self._plan is None:
            self.get_plan(state)
        self._replan_counter += 1
        if REPLAN_IF_FAILURE and self._need_replan and self._replan_counter % RETRY_FOR_NEW_PLAN_EVERY == 0:
            self.get_plan(state)
        block, goal = self._goal_and_block(state)
        while len(self._plan) > 1 and np.linalg.norm(self._plan[0] - block) <= ADVANCE_TO_NEXT_SUBGOAL_THRESHOLD:
            self._plan.popleft()
            self._ee_plan = None                              # new push direction: new pre-push pose
        waypoint = self._plan[0]

        free_space = False
        if self._use_ee_planner and self.phase == "move_to_pre_block":
            delta = self._free_space_action(state, block, waypoint)
            if delta is not None:
                n = np.linalg.norm(delta)
                max_step = self._max_step
                if self._slowdown_freespace:
                    max_step *= slowdown_factor(float(np.linalg.norm(goal - block)))
                act = (delta / n * min(n, max_step)) if n > 1e-9 else np.zeros(2)
                free_space = True
        if not free_space:
            # aim the oriented-push state machine at the current block subgoal
            saved = self._goal_and_block
            self._goal_and_block = lambda st: (block, waypoint)
            try:
                act = super().action(state)
            finally:
                self._goal_and_block = saved
            act = np.asarray(act, float) * slowdown_factor(float(np.linalg.norm(goal - block)))
            if self.phase == "move_to_pre_block":
                self._ee_plan = None                          # lost the block: the next free-space phase plans afresh
        if free_space and self._noise > 0:
            act = act + self._rng.normal(0, self._noise, 2)
        return np.asarray(act, np.float32)
