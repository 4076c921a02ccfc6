"""Run with the REFERENCE's ``language_table`` first on sys.path: reads a state file this repository wrote with the reference's
own ``read_pybullet_state``, reports what it found, and writes the records back out with the reference's ``write_pybullet_state``
for this repository to read.  Driven by tests/test_reference_parity_data.py."""
import json
import sys

from language_table.environments.utils import utils_pybullet as up


def main(src, dst, report):
    data = up.read_pybullet_state(src)
    st = data["pybullet_state"]
    robot = st["robots"][0]
    assert isinstance(robot, up.XarmState) and all(isinstance(o, up.ObjState) for o in st["objects"])
    out = {"task": data["task"], "version": data["state_version"], "n_objects": len(st["objects"]),
           "object_poses": [[list(o.base_pose[0]), list(o.base_pose[1])] for o in st["objects"]],
           "joint_positions": [js[0] for js in robot.joint_state], "joint_names": [ji[1] for ji in robot.joint_info],
           "target_translation": [float(v) for v in robot.target_effector_pose.translation],
           "instruction": st["instruction"], "actions": data["actions"]}
    json.dump(out, open(report, "w"))
    up.write_pybullet_state(dst, st, data["task"], data["actions"])


if __name__ == "__main__":
    main(*sys.argv[1:4])
