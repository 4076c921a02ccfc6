"""Outputs of the camera-ray helpers of ``language_table.environments.utils.utils_pybullet`` (whichever package is first on
sys.path) on fixed random inputs: matrix layout conversion, view-matrix decomposition, pixel → view ray → world ray,
ray / plane intersection.  tests/test_reference_parity_data.py compares the reference's numbers with this repository's."""
import json
import sys

import numpy as np
from scipy.spatial import transform

from language_table.environments.utils import utils_pybullet as up


def main(path):
    rng = np.random.RandomState(0)
    rows = []
    for _ in range(8):
        rot = transform.Rotation.from_rotvec(rng.uniform(-1.5, 1.5, 3)).as_matrix()
        eye = rng.uniform(-1, 1, 3)
        view = np.eye(4)
        view[:3, :3], view[:3, 3] = rot, -rot @ eye
        view_gl = view.T.reshape(-1).tolist()                              # column-major, the order Bullet / OpenGL hand out
        fov, aspect, near, far = rng.uniform(0.6, 1.2), 320.0 / 180.0, 0.01, 10.0
        f = 1.0 / np.tan(fov / 2)
        proj = np.array([[f / aspect, 0, 0, 0], [0, f, 0, 0], [0, 0, (far + near) / (near - far), 2 * far * near / (near - far)],
                         [0, 0, -1, 0]])
        try:
            xyz, quat = up.decompose_view_matrix(view_gl)
            quat = np.asarray(quat) * (1.0 if quat[3] >= 0 else -1.0)      # q and -q are the same rotation
        except ValueError:         # the reference hands a 4x4 to Rotation.from_matrix, which current scipy rejects
            xyz, quat = np.linalg.inv(view)[:3, 3], np.full(4, np.nan)
        px = rng.uniform([0, 0], [320, 180])
        o, d = up.image_xy_to_view_ray(px, 320, 180, np.linalg.inv(proj))
        wo, wd = up.view_ray_to_world_ray(o, d, np.linalg.inv(view))
        hit = up.ray_to_plane_test(np.asarray(wo), np.asarray(wd), np.array([0.35, 0.0, 0.0]), np.array([0.0, 0.0, 1.0]))
        miss = up.ray_to_plane_test(np.array([0.0, 0.0, 1.0]), np.array([1.0, 0.0, 0.0]), np.zeros(3), np.array([0.0, 0.0, 1.0]))
        behind = up.ray_to_plane_test(np.array([0.0, 0.0, 1.0]), np.array([0.0, 0.0, 1.0]), np.zeros(3), np.array([0.0, 0.0, 1.0]))
        rows.append({"mat": np.round(up.pybullet_mat_to_numpy_4x4(view_gl), 9).tolist(), "xyz": np.round(xyz, 9).tolist(),
                     "quat": None if np.isnan(quat).any() else np.round(quat, 9).tolist(), "view_ray": np.round(np.concatenate([o, d]), 9).tolist(),
                     "world_ray": np.round(np.concatenate([wo, wd]), 9).tolist(),
                     "hit": None if hit is None else np.round(hit, 7).tolist(), "miss": miss is None, "behind": behind is None})
    r = transform.Rotation.from_rotvec([0.3, -0.2, 0.9])
    m = up.rotation_to_matrix(r)
    rows.append({"rotation_to_matrix": np.round(np.asarray(m), 9).tolist(),
                 "matrix_to_rotation": np.round(up.matrix_to_rotation(np.asarray(m)).as_quat(), 9).tolist()})
    json.dump(rows, open(path, "w"))


if __name__ == "__main__":
    main(sys.argv[1])
