"""Plans the same 2-D routing problems with whichever ``language_table`` package is first on sys.path, through the public
entry point ``rrt_star.get_shortest_path_no_collisions`` with the push oracle's planner settings, and records success, path
length, clearance and wall time per problem.  tests/test_reference_parity_data.py compares the reference's record with ours."""
import json
import sys
import time

import numpy as np

from language_table.environments.oracles import rrt_star


def problems(n=12):
    rng = np.random.RandomState(0)
    out = []
    while len(out) < n:
        start, goal = rng.uniform([0.2, -0.25], [0.55, 0.25]), rng.uniform([0.2, -0.25], [0.55, 0.25])
        if np.linalg.norm(goal - start) < 0.2:
            continue
        obs = [o for o in rng.uniform([0.18, -0.28], [0.58, 0.28], (7, 2))
               if np.linalg.norm(o - start) > 0.06 and np.linalg.norm(o - goal) > 0.06]
        out.append((start, goal, np.asarray(obs)))
    return out


def main(path):
    rows = []
    np.random.seed(0)                                          # the reference's planner draws from numpy's global generator
    for start, goal, obs in problems():
        t0 = time.perf_counter()
        route, ok = rrt_star.get_shortest_path_no_collisions(
            tuple(start), tuple(goal), (0.15, 0.64), (-0.3448, 0.3448), [tuple(o) for o in obs], [0.015] * len(obs),
            delta=0.015, step_length=0.05, goal_sample_rate=0.1, search_radius=0.5, iter_max=1024)
        dt = time.perf_counter() - t0
        pts = np.asarray([np.asarray(p, float) for p in route])
        seg = np.linalg.norm(np.diff(pts, axis=0), axis=1)
        clearance = min((float(np.min(np.linalg.norm(pts[:, None] - obs[None], axis=-1))) if len(obs) else 1.0), 1.0)
        rows.append({"ok": bool(ok), "length": float(seg.sum()), "straight": float(np.linalg.norm(goal - start)),
                     "first_is_goal": bool(np.allclose(pts[0], goal)), "last_is_start": bool(np.allclose(pts[-1], start)),
                     "waypoint_clearance": clearance, "seconds": dt})
    json.dump(rows, open(path, "w"))


if __name__ == "__main__":
    main(sys.argv[1])
