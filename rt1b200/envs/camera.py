"""Camera matrices of the Language-Table scene in the OpenGL convention PyBullet uses, computed without PyBullet.

``calc_camera_params`` mirrors what the reference obtains from ``computeViewMatrix`` / ``computeProjectionMatrixFOV``
(``language_table/environments/language_table.py:365-397``): both matrices as 16-element column-major tuples, plus the
camera position, look-at point and up vector.  The helpers below convert them to numpy and project points; the ray
helpers live in ``language_table/environments/utils/utils_pybullet.py`` under the reference's names.
"""
from __future__ import annotations

import math
from typing import Sequence, Tuple

import numpy as np


def quaternion_from_euler(rpy: Sequence[float]) -> np.ndarray:
    """Bullet's getQuaternionFromEuler: intrinsic z-y'-x'' (yaw, pitch, roll) → (x, y, z, w)."""
    r, p, y = (0.5 * a for a in rpy)
    cr, sr, cp, sp, cy, sy = math.cos(r), math.sin(r), math.cos(p), math.sin(p), math.cos(y), math.sin(y)
    return np.array([sr * cp * cy - cr * sp * sy, cr * sp * cy + sr * cp * sy, cr * cp * sy - sr * sp * cy,
                     cr * cp * cy + sr * sp * sy])


def matrix_from_quaternion(q: Sequence[float]) -> np.ndarray:
    x, y, z, w = q
    return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                     [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                     [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])


def quaternion_from_matrix(m: np.ndarray) -> np.ndarray:
    """Rotation matrix → unit quaternion (x, y, z, w), numerically stable branch selection."""
    m = np.asarray(m, np.float64)
    t = np.trace(m)
    if t > 0:
        s = math.sqrt(t + 1.0) * 2
        q = [(m[2, 1] - m[1, 2]) / s, (m[0, 2] - m[2, 0]) / s, (m[1, 0] - m[0, 1]) / s, 0.25 * s]
    else:
        i = int(np.argmax(np.diag(m)))
        j, k = (i + 1) % 3, (i + 2) % 3
        s = math.sqrt(max(1.0 + m[i, i] - m[j, j] - m[k, k], 1e-12)) * 2
        q = [0.0, 0.0, 0.0, (m[k, j] - m[j, k]) / s]
        q[i], q[j], q[k] = 0.25 * s, (m[j, i] + m[i, j]) / s, (m[k, i] + m[i, k]) / s
    q = np.asarray(q)
    return q / np.linalg.norm(q)


def compute_view_matrix(eye, target, up) -> Tuple[float, ...]:
    """Right-handed look-at matrix, 16 values column-major (what pybullet.computeViewMatrix returns)."""
    eye, target, up = (np.asarray(v, np.float64) for v in (eye, target, up))
    f = target - eye
    f /= np.linalg.norm(f)
    s = np.cross(f, up / np.linalg.norm(up))
    s /= np.linalg.norm(s)
    u = np.cross(s, f)
    m = np.eye(4)
    m[0, :3], m[1, :3], m[2, :3] = s, u, -f
    m[:3, 3] = -m[:3, :3] @ eye
    return tuple(m.T.reshape(-1))


def compute_projection_matrix_fov(fov_deg: float, aspect: float, near: float, far: float) -> Tuple[float, ...]:
    """OpenGL perspective projection from the VERTICAL field of view, column-major."""
    f = 1.0 / math.tan(math.radians(fov_deg) / 2)
    m = np.zeros((4, 4))
    m[0, 0], m[1, 1] = f / aspect, f
    m[2, 2], m[2, 3] = (far + near) / (near - far), 2 * far * near / (near - far)
    m[3, 2] = -1.0
    return tuple(m.T.reshape(-1))


def calc_camera_params(image_size, camera_pose, camera_orientation, intrinsics, zrange=(0.01, 10.0)):
    """``image_size = (height, width)`` → (view matrix, projection matrix, position, look-at, up)."""
    rot = matrix_from_quaternion(quaternion_from_euler(camera_orientation))
    lookdir = rot @ np.array([0.0, 0.0, 1.0])
    updir = rot @ np.array([0.0, -1.0, 0.0])
    position = np.asarray(camera_pose, np.float64)
    lookat = position + lookdir
    fov = math.degrees(2 * math.atan((image_size[0] / 2) / intrinsics[0]))
    viewm = compute_view_matrix(position, lookat, updir)
    projm = compute_projection_matrix_fov(fov, image_size[1] / image_size[0], zrange[0], zrange[1])
    return viewm, projm, position, lookat, updir


def mat4(column_major: Sequence[float]) -> np.ndarray:
    assert len(column_major) == 16, "a 4x4 matrix has 16 elements"
    return np.asarray(column_major, np.float64).reshape(4, 4).T


def project_points(points_world: np.ndarray, viewm, projm, width: int, height: int) -> np.ndarray:
    """World points ``[..., 3]`` → pixel coordinates ``[..., 2]`` (x right, y down) through view and projection."""
    p = np.concatenate([np.asarray(points_world, np.float64), np.ones(points_world.shape[:-1] + (1,))], -1)
    clip = p @ (mat4(projm) @ mat4(viewm)).T
    ndc = clip[..., :3] / clip[..., 3:4]
    return np.stack([(ndc[..., 0] + 1) * 0.5 * width, (1 - ndc[..., 1]) * 0.5 * height], -1)
