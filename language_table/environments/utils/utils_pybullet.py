"""State (de)serialisation helpers (reference ``environments/utils/utils_pybullet.py:252-446``): object/arm state
records and gz-compressed JSON state files.  The URDF loading / Bullet ray-cast helpers of the original have no
counterpart (no physics engine on this box)."""
import dataclasses
import gzip
import json
from typing import Any, Dict, List, Optional

import numpy as np


@dataclasses.dataclass
class ObjState:
    obj_id: int
    base_pose: Any
    base_vel: Any = ((0, 0, 0), (0, 0, 0))
    joint_info: Any = ()
    joint_state: Any = ()

    def serialize(self) -> Dict[str, Any]:
        return {"obj_id": self.obj_id, "base_pose": _tolist(self.base_pose), "base_vel": _tolist(self.base_vel),
                "joint_info": _tolist(self.joint_info), "joint_state": _tolist(self.joint_state)}

    @staticmethod
    def deserialize(data: Dict[str, Any]) -> "ObjState":
        return ObjState(data["obj_id"], data["base_pose"], data.get("base_vel"), data.get("joint_info", ()), data.get("joint_state", ()))


@dataclasses.dataclass
class XarmState(ObjState):
    target_effector_pose: Optional[Any] = None
    goal_translation: Optional[Any] = None

    def serialize(self) -> Dict[str, Any]:
        d = super().serialize()
        d["target_effector_pose"] = None if self.target_effector_pose is None else self.target_effector_pose.serialize()
        d["goal_translation"] = None if self.goal_translation is None else _tolist(self.goal_translation)
        return d


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
