"""Camera-ray geometry and state (de)serialisation helpers of the reference's ``environments/utils/utils_pybullet.py``.

* view-matrix decomposition, pixel → view ray → world ray, ray/plane intersection and the workspace plane
  (reference ``:115-206``) — plain linear algebra, no Bullet client involved;
* object/arm state records and gz-compressed JSON state files (reference ``:252-446``).
``load_urdf`` of the original is replaced by the parsers in ``rt1b200.envs.assets`` (no physics engine on this box).
"""
import dataclasses
import gzip
import json
from typing import Any, Dict, List, Optional, Tuple

import numpy as np

Vec3 = Tuple[float, float, float]
Vec4 = Tuple[float, float, float, float]

from rt1b200.envs import camera as _cam


def pybullet_mat_to_numpy_4x4(pybullet_matrix):
    """16-element column-major matrix (Bullet / OpenGL order) → 4x4 numpy array."""
    return _cam.mat4(pybullet_matrix)


def decompose_view_matrix(pybullet_view_matrix):
    """Rigid view matrix → (camera position in the world, camera orientation as an (x, y, z, w) quaternion)."""
    view_to_world = np.linalg.inv(pybullet_mat_to_numpy_4x4(pybullet_view_matrix))
    return view_to_world[:3, 3].copy(), _cam.quaternion_from_matrix(view_to_world[:3, :3])


def world_obj_to_view(world_xyz_obj, world_quat_obj, camera_view, client=None):
    """Pose of an object expressed in the camera (view) frame; ``client`` is accepted for signature parity and unused."""
    world_to_view = pybullet_mat_to_numpy_4x4(camera_view)
    xyz = world_to_view[:3, :3] @ np.asarray(world_xyz_obj, np.float64) + world_to_view[:3, 3]
    rot = world_to_view[:3, :3] @ _cam.matrix_from_quaternion(world_quat_obj)
    return xyz, _cam.quaternion_from_matrix(rot)


def image_xy_to_view_ray(xy, cam_width, cam_height, proj_mat_inv):
    """Pixel (x right, y down) → ray through it in view space: (origin = camera centre, unit direction)."""
    ndc = np.array([2.0 * xy[0] / cam_width - 1.0, 1.0 - 2.0 * xy[1] / cam_height, 0.0, 1.0])
    eye = np.asarray(proj_mat_inv, np.float64) @ ndc
    d = eye[:3]
    return np.zeros(3), d / max(float(np.linalg.norm(d)), 1e-6)


def view_ray_to_world_ray(origin, vec, view_mat_inv):
    m = np.asarray(view_mat_inv, np.float64)
    return (m @ np.append(origin, 1.0))[:3], (m @ np.append(vec, 0.0))[:3]


def ray_to_plane_test(ray_origin, ray_vec, plane_origin, plane_normal):
    """Intersection point of a ray with a plane, or ``None`` (parallel, or the plane is behind the origin)."""
    ray_origin, ray_vec = np.asarray(ray_origin, np.float64), np.asarray(ray_vec, np.float64)
    plane_origin, plane_normal = np.asarray(plane_origin, np.float64), np.asarray(plane_normal, np.float64)
    denom = float(plane_normal @ ray_vec)
    if abs(denom) < np.finfo(np.float32).eps:
        return None
    t = float(plane_normal @ (plane_origin - ray_origin)) / denom
    return ray_origin + t * ray_vec if t >= 0 else None


def get_workspace(env=None):
    """Origin and normal of the board plane (the reference queries Bullet for the pose of ``workspace_real.urdf``,
    which it spawns at (0.35, 0, 0) with identity orientation: ``language_table.py:560-563``)."""
    return np.array([0.35, 0.0, 0.0]), np.array([0.0, 0.0, 1.0])


def pixel_to_workspace(xy, viewm, projm, cam_width, cam_height):
    """Where the ray through pixel ``xy`` meets the board — the composition the reference's GUI tools perform."""
    o, d = image_xy_to_view_ray(xy, cam_width, cam_height, np.linalg.inv(pybullet_mat_to_numpy_4x4(projm)))
    o, d = view_ray_to_world_ray(o, d, np.linalg.inv(pybullet_mat_to_numpy_4x4(viewm)))
    return ray_to_plane_test(o, d, *get_workspace())


def _lists_to_tuple(obj):
    return tuple(_lists_to_tuple(v) for v in obj) if isinstance(obj, (list, tuple)) else obj


@dataclasses.dataclass
class ObjState:
    """One body of the scene: base pose ((x, y, z), quaternion), base velocity, and per joint its info tuple and state
    tuple — read from / written to a Bullet-style client (reference ``utils_pybullet.py:252-332``;
    ``board_client.BoardClient`` is the client of this environment, a real ``pybullet`` client works as well)."""
    obj_id: int
    base_pose: Any
    base_vel: Any = ((0, 0, 0), (0, 0, 0))
    joint_info: Any = ()
    joint_state: Any = ()

    @staticmethod
    def _get_joint_info(client, obj_id, joint_index):
        return tuple(v.decode("utf-8") if isinstance(v, bytes) else v for v in client.getJointInfo(obj_id, joint_index))

    @staticmethod
    def get_bullet_state(client, obj_id) -> "ObjState":
        n = client.getNumJoints(obj_id)
        return ObjState(obj_id=obj_id, base_pose=client.getBasePositionAndOrientation(obj_id),
                        base_vel=client.getBaseVelocity(obj_id),
                        joint_info=tuple(ObjState._get_joint_info(client, obj_id, i) for i in range(n)),
                        joint_state=tuple(client.getJointState(obj_id, i) for i in range(n)))

    def set_bullet_state(self, client, obj_id) -> None:
        """Hard-set the body to this record.  The joints must be the ones the record was taken from: count, order and
        info tuples are checked before anything is written to a joint."""
        xyz, quat = self.base_pose
        client.resetBasePositionAndOrientation(obj_id, xyz, quat)
        vel, ang_vel = self.base_vel if self.base_vel is not None else ((0, 0, 0), (0, 0, 0))
        client.resetBaseVelocity(obj_id, vel, ang_vel)
        n = client.getNumJoints(obj_id)
        if n != len(self.joint_info) or n != len(self.joint_state):
            raise ValueError("Incorrect number of joint info state pairs.")
        for i, (info, state) in enumerate(zip(self.joint_info, self.joint_state)):
            if info[0] != i:
                raise ValueError("Joint index mismatch.")
            current = ObjState._get_joint_info(client, obj_id, i)
            if _lists_to_tuple(current) != _lists_to_tuple(info):
                raise ValueError("joint_info mismatch %s vs %s (expected)" % (str(current), str(info)))
        for i, state in enumerate(self.joint_state):
            client.resetJointState(obj_id, i, targetValue=state[0], targetVelocity=state[1])

    def serialize(self) -> Dict[str, Any]:
        return {"obj_id": self.obj_id, "base_pose": _tolist(self.base_pose), "base_vel": _tolist(self.base_vel),
                "joint_info": _tolist(self.joint_info), "joint_state": _tolist(self.joint_state)}

    @staticmethod
    def deserialize(data: Dict[str, Any]) -> "ObjState":
        return ObjState(data["obj_id"], _lists_to_tuple(data["base_pose"]), _lists_to_tuple(data.get("base_vel")),
                        _lists_to_tuple(data.get("joint_info", ())), _lists_to_tuple(data.get("joint_state", ())))


@dataclasses.dataclass
class XarmState(ObjState):
    """The arm: an ``ObjState`` plus the controller's set point and the goal marker (reference ``:335-373``)."""
    target_effector_pose: Optional[Any] = None
    goal_translation: Optional[Any] = None

    @staticmethod
    def get_bullet_state(client, obj_id, target_effector_pose=None, goal_translation=None) -> "XarmState":
        if goal_translation is not None:
            goal_translation = tuple(np.asarray(goal_translation).tolist())
        base = ObjState.get_bullet_state(client, obj_id)
        return XarmState(obj_id=base.obj_id, base_pose=base.base_pose, base_vel=base.base_vel, joint_info=base.joint_info,
                         joint_state=base.joint_state, target_effector_pose=target_effector_pose,
                         goal_translation=goal_translation)

    def serialize(self) -> Dict[str, Any]:
        d = super().serialize()
        d["target_effector_pose"] = None if self.target_effector_pose is None else self.target_effector_pose.serialize()
        d["goal_translation"] = [] if self.goal_translation is None else _tolist(self.goal_translation)
        return d

    @staticmethod
    def deserialize(data: Dict[str, Any]) -> "XarmState":
        from language_table.environments.utils.pose3d import Pose3d
        tep, goal = data.get("target_effector_pose"), data.get("goal_translation")
        base = ObjState.deserialize(data)
        return XarmState(obj_id=base.obj_id, base_pose=base.base_pose, base_vel=base.base_vel, joint_info=base.joint_info,
                         joint_state=base.joint_state, target_effector_pose=None if tep is None else Pose3d.deserialize(tep),
                         goal_translation=_lists_to_tuple(goal) if goal else None)


def _tolist(x):
    if isinstance(x, np.ndarray):
        return x.tolist()
    if isinstance(x, (list, tuple)):
        return [_tolist(v) for v in x]
    if hasattr(x, "tolist"):
        return x.tolist()
    return x


def write_state_file(path: str, states: List[Dict[str, Any]]) -> None:
    with gzip.open(path, "wt") as f:
        json.dump([{k: _tolist(v) for k, v in s.items() if k not in ("reward", "task_info")} for s in states], f)


def read_state_file(path: str) -> List[Dict[str, Any]]:
    with gzip.open(path, "rt") as f:
        return json.load(f)


# ------------------------------------------------------------------------------ reference-format state files (version 2)
PYBULLET_STATE_VERSION = 2


def rotation_to_matrix(rotation):
    return rotation.as_matrix()


def matrix_to_rotation(matrix):
    from scipy.spatial import transform
    return transform.Rotation.from_matrix(matrix)


def _serialize_pybullet_state(state):
    """Plain-old-data form of a state dict whose leaves are ints or ObjState / XarmState records (reference ``:376-393``)."""
    if isinstance(state, list):
        return [_serialize_pybullet_state(e) for e in state]
    if isinstance(state, dict):
        assert "_serialized_obj_name" not in state
        return {k: _serialize_pybullet_state(v) for k, v in state.items()}
    if isinstance(state, (XarmState, ObjState)):
        return {"_serialized_obj_name": type(state).__name__, "_serialized_data": state.serialize()}
    if isinstance(state, (int, float, str)) or state is None:
        return state
    raise ValueError("Unhandled type for object %s, type %s" % (str(state), type(state)))


def _deserialize_pybullet_state(state):
    if isinstance(state, list):
        return [_deserialize_pybullet_state(e) for e in state]
    if isinstance(state, dict):
        if "_serialized_obj_name" in state:
            name, data = state["_serialized_obj_name"], state["_serialized_data"]
            if name == XarmState.__name__:
                return XarmState.deserialize(data)
            if name == ObjState.__name__:
                return ObjState.deserialize(data)
            raise ValueError("Unsupported: %s" % name)
        return {k: _deserialize_pybullet_state(v) for k, v in state.items()}
    return state


def write_pybullet_state(filename: str, pybullet_state, task, actions=None) -> None:
    """gz-compressed JSON in the reference's layout (``:417-430``): ``pybullet_state`` (dict of ObjState / XarmState lists),
    ``state_version``, ``ts_ms``, ``user``, ``task``, ``actions``."""
    import getpass
    import time
    data = {"pybullet_state": _serialize_pybullet_state(pybullet_state), "state_version": PYBULLET_STATE_VERSION,
            "ts_ms": int(time.time()) * 1000, "user": getpass.getuser(), "task": task,
            "actions": [] if actions is None else _tolist(actions)}
    with gzip.open(filename, "wb") as f:
        f.write(json.dumps(data).encode("utf-8"))


def read_pybullet_state(filename: str):
    with gzip.open(filename, "rb") as f:
        data = json.loads(f.read().decode("utf-8"))
    assert isinstance(data, dict)
    if data["state_version"] != PYBULLET_STATE_VERSION:
        raise ValueError("incompatible state data (version %d, expected %d)" % (data["state_version"], PYBULLET_STATE_VERSION))
    data["pybullet_state"] = _deserialize_pybullet_state(data["pybullet_state"])
    return data


# ------------------------------------------------------------------------------------------------ assets without Bullet
def resolve_asset_path(file_path: str) -> str:
    """A URDF / OBJ path in the reference's convention → a file on this machine.  The reference's ``load_urdf`` (``:61-91``)
    tries the path as given, then strips ``third_party/py/`` and resolves ``language_table/environments/...`` against the
    package root (Bullet's own data under ``third_party/bullet/...`` comes from ``pybullet_data``, which does not exist
    here).  The same rules, against the asset directory ``rt1b200.envs.assets.find_asset_dir()`` found."""
    import os
    from rt1b200.envs import assets
    if os.path.exists(file_path):
        return file_path
    rel = file_path[len("third_party/py/"):] if file_path.startswith("third_party/py/language_table/") else file_path
    marker = "language_table/environments/assets/"
    root = assets.find_asset_dir()
    if marker in rel and root is not None:
        cand = os.path.join(root, rel.split(marker, 1)[1])
        if os.path.exists(cand):
            return cand
    raise FileNotFoundError("Cannot load the URDF file {}".format(file_path))


def load_urdf(pybullet_client, file_path, *args, **kwargs):
    """Without a physics engine "loading" a URDF means parsing its visuals: returns the list of
    ``{"tris": [T,3,3], "rgb": (r,g,b)}`` records of :func:`rt1b200.envs.assets.load_urdf_visuals` (``pybullet_client`` and
    Bullet's placement arguments are accepted and ignored)."""
    from rt1b200.envs import assets
    return assets.load_urdf_visuals(resolve_asset_path(file_path))


def add_visual_sphere(client=None, center=(0, 0, 0), radius=0.1, rgba=(0.5, 0.5, 0.5, 0.5)):
    """GUI debugging aid of the reference (a non-colliding sphere in the Bullet visualiser): there is no visualiser on this
    box, so this only returns the description of what would have been drawn."""
    return {"center": tuple(center), "radius": float(radius), "rgba": tuple(rgba)}


def reset_camera_pose(env, view_type):
    """The reference re-aims the Bullet debug visualiser's camera (``:208-241``).  No visualiser here; returns the
    ``(distance, yaw_deg, pitch_deg, target)`` it would have set, computed with the same rules."""
    if view_type == "POLICY":
        viewm, _, position, lookat, _ = env.calc_camera_params((180, 320))
        euler = matrix_to_rotation(pybullet_mat_to_numpy_4x4(viewm)[0:3, 0:3]).as_euler("xyz", degrees=False)
        pitch, yaw = euler[1], -euler[2]
        distance = float(np.linalg.norm(np.asarray(position) - np.asarray(lookat)) * 0.6)
        target = np.asarray(lookat)
    elif view_type == "TOP_DOWN":
        target, _ = get_workspace(env)
        distance, yaw, pitch = 0.5, np.pi / 2, -(np.pi / 2 - 1e-5)
    else:
        raise ValueError("unsupported view_type %s" % view_type)
    return distance, 360 * yaw / (2.0 * np.pi), 360 * pitch / (2.0 * np.pi), target
