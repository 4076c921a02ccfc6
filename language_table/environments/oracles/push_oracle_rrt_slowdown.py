"""Obstacle-aware pushing oracle (reference ``environments/oracles/push_oracle_rrt_slowdown.py:95``): plan the block's
route to the goal with RRT* around the other blocks, push it waypoint by waypoint with the oriented-push state machine,
and slow down near the goal.  ``get_plan`` is what the evaluator uses to reject unsolvable initial boards
(``eval/main_rt1.py:162-172``)."""
import numpy as np

from language_table.environments import constants
from language_table.environments.oracles import oriented_push_oracle, rrt_star


class ObstacleOrientedPushOracleBoard2dRRT(oriented_push_oracle.OrientedPushOracle):
    def __init__(self, env, use_ee_planner=True, action_noise_std=0.0, max_step=0.03, slowdown_radius=0.08, rng=None):
        super().__init__(env, action_noise_std=action_noise_std, max_step=max_step, rng=rng)
        self._use_ee_planner = use_ee_planner
        self._slowdown_radius = slowdown_radius
        self._plan = None
        self._waypoint = 0

    def reset(self):
        super().reset()
        self._plan, self._waypoint = None, 0

    def _obstacles(self, state, exclude):
        obs = []
        for k, v in state.items():
            if k.startswith("block_") and k.endswith("_translation"):
                name = k[len("block_"):-len("_translation")]
                if name in exclude or float(state.get("block_%s_mask" % name, [1.0])[0]) == 0.0:
                    continue
                obs.append((float(v[0]), float(v[1]), 2 * constants.BLOCK_RADIUS + 0.005))
        return obs

    def get_plan(self, state):
        """Plan the pushed block's path; returns True when a collision-free route exists (or none is needed)."""
        info = getattr(self._env, "_task_info", None)
        block = getattr(info, "block1", None) or getattr(info, "block", None) or getattr(info, "block_target", None)
        if block is None:
            return False
        exclude = {block}
        if hasattr(info, "block2"):
            exclude.add(info.block2)
        start, goal = self._goal_and_block(state)
        bounds = ((constants.X_MIN, constants.Y_MIN), (constants.X_MAX, constants.Y_MAX))
        planner = rrt_star.RRTStar(start, goal, self._obstacles(state, exclude), bounds, rng=self._rng, max_iter=1024)
        path = planner.plan()
        if path is None:
            planner = rrt_star.RRTStar(start, goal, self._obstacles(state, exclude), bounds, rng=self._rng, max_iter=2048,
                                       expand_dis=0.02)
            path = planner.plan()
        self._plan, self._waypoint = path, 1
        return path is not None

    def action(self, state=None):
        state = state or self._env.compute_state()
        if self._plan is None and not self.get_plan(state):
            return super().action(state)                                    # potential-field style fallback: push straight
        block, goal = self._goal_and_block(state)
        while self._waypoint < len(self._plan) - 1 and np.linalg.norm(self._plan[self._waypoint] - block) < 0.03:
            self._waypoint += 1
        # temporarily aim the oriented-push machine at the current waypoint
        waypoint = self._plan[min(self._waypoint, len(self._plan) - 1)]
        saved = self._goal_and_block
        self._goal_and_block = lambda st: (block, waypoint)
        try:
            act = super().action(state)
        finally:
            self._goal_and_block = saved
        if np.linalg.norm(goal - block) < self._slowdown_radius:
            act = act * 0.5
        return act
