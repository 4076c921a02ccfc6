"""``LanguageTable``: the single-board environment with the reference's interface
(``language_table/environments/language_table.py:45``: ``reset / step / render / succeeded / compute_state /
get_pybullet_state / set_pybullet_state``), on top of the PyBullet-free board in :mod:`rt1b200.envs.board`.

Observation (``:407-416``): ``effector_translation f32[2]``, ``effector_target_translation f32[2]``,
``instruction int32[512]`` (zero-padded utf-8 bytes), ``rgb uint8[180,320,3]``.  ``compute_state`` adds per-block
``block_<name>_translation / _orientation / _mask`` entries used by the rewards and oracles.  Rendering is the
z-buffer rasteriser of :mod:`rt1b200.envs.raster` (CUDA when the board lives on a GPU, numpy otherwise); PyBullet's
TinyRenderer / GUI paths are unavailable on this box and have no counterpart here.
"""
from __future__ import annotations

import collections
import copy
from typing import Any, Dict, Optional

import numpy as np
import torch

from language_table.environments import blocks as blocks_module
from language_table.environments import constants
from language_table.environments.rewards import task_info
from language_table.environments.utils import xarm_sim_robot
from language_table.environments.utils.pose3d import Pose3d

from rt1b200 import spaces
from rt1b200.envs.board import BatchedLanguageTable


def add_debug_info_to_image(image: np.ndarray, info_dict: Dict[str, Any], pos=(0, 0), text_color=(0, 0, 0)) -> np.ndarray:
    """The frame scaled to 640x360 under a white band that carries ``instruction: <text>``, wrapped at 35 characters
    (reference ``language_table.py:1000-1030``, which draws with OpenCV's Hershey font; OpenCV is not available offline, so
    the text is set with Pillow's built-in font — same layout, different glyphs).  Returns uint8 ``[444, 640, 3]``."""
    import textwrap
    from PIL import Image, ImageDraw, ImageFont
    img = Image.fromarray(np.ascontiguousarray(image, dtype=np.uint8)).resize((640, 360), Image.BILINEAR)
    band = int(3 * int(360 * 0.08))
    canvas = Image.new("RGB", (640, 360 + band), (255, 255, 255))
    canvas.paste(img, (0, band))
    text = "instruction: %s" % info_dict["instruction"] if "instruction" in info_dict else ""
    try:
        font = ImageFont.load_default(size=24)           # Pillow >= 10.1 with FreeType
    except Exception:                                     # noqa: BLE001 - older Pillow / no FreeType: the small bitmap font
        font = ImageFont.load_default()
    draw = ImageDraw.Draw(canvas)
    x, y = pos
    for line in textwrap.wrap(text, width=35):
        box = draw.textbbox((x, y), line, font=font)
        draw.text((x, y), line, fill=tuple(text_color), font=font)
        y += int((box[3] - box[1]) * 1.2) + 2
    return np.asarray(canvas, dtype=np.uint8)


def sleep_spin(sleep_time_sec: float) -> None:
    """Busy-wait sleep (reference ``:1032-1038``; used when pacing a viewer to the control rate)."""
    import time
    if sleep_time_sec <= 0:
        return
    t0 = time.perf_counter()
    while time.perf_counter() - t0 < sleep_time_sec:
        pass


class LanguageTable:
    metadata = {"render.modes": ["rgb_array"]}

    def __init__(self, block_mode, reward_factory=None, control_frequency=10.0, shared_memory=False, seed=None,
                 delay_reward_steps=0, render_text_in_image=True, training=True, device="cpu", arm_kinematics=True,
                 debug_visuals=False, add_invisible_walls=False):
        if shared_memory:
            raise NotImplementedError("the PyBullet GUI / shared-memory visualiser is unavailable on this box")
        self._block_mode = blocks_module.LanguageTableBlockVariants(block_mode) if isinstance(block_mode, str) else block_mode
        self._training = training
        self._rng = np.random.RandomState(seed=seed)
        self._control_frequency = control_frequency
        self._step_frequency = constants.STEP_FREQUENCY
        self._sim_steps_per_step = int(self._step_frequency / self._control_frequency)     # 24, reference :114-124
        self._render_text_in_image = render_text_in_image
        # the xArm6 that carries the cylinder: IK to the commanded pose, joint servo over the simulation sub-steps, so the
        # effector lags its target within a control step like the motor-driven arm of the reference (:599-646)
        self._robot = xarm_sim_robot.XArmSimRobot(initial_joint_positions=constants.INITIAL_JOINT_POSITIONS,
                                                  end_effector="cylinder_real", color="white") if arm_kinematics else None
        self._board = BatchedLanguageTable(1, self._block_mode if self._block_mode in (
            blocks_module.LanguageTableBlockVariants.BLOCK_1, blocks_module.LanguageTableBlockVariants.BLOCK_4,
            blocks_module.LanguageTableBlockVariants.BLOCK_8, blocks_module.LanguageTableBlockVariants.N_CHOOSE_K)
            else blocks_module.LanguageTableBlockVariants.BLOCK_8, device=device, seed=seed or 0, max_episode_steps=10 ** 9,
            add_invisible_walls=add_invisible_walls)
        self._debug_visuals = debug_visuals        # GUI-only in the reference (debug lines in the PyBullet visualiser): accepted, no effect
        self._all_blocks = self._board.block_names
        self._blocks_on_table = list(self._all_blocks)
        self._reward_calculator = None
        if reward_factory is not None:
            self._reward_calculator = reward_factory(goal_reward=100.0, rng=self._rng, delay_reward_steps=delay_reward_steps,
                                                     block_mode=self._block_mode)
        self._instruction = np.zeros(constants.INSTRUCTION_LENGTH, np.int32)
        self._instruction_str = None
        self._task_info = None
        self._start_block = self._oracle_target_block = self._oracle_target_translation = None
        self._target_absolute_location = self._target_relative_location = None
        self.action_space = spaces.Box(low=-0.1, high=0.1, shape=(2,), dtype=np.float32)
        self.observation_space = spaces.Dict(collections.OrderedDict(
            effector_translation=spaces.Box(low=constants.WORKSPACE_BOUNDS[0] - 0.1, high=constants.WORKSPACE_BOUNDS[1] + 0.1),
            effector_target_translation=spaces.Box(low=constants.WORKSPACE_BOUNDS[0] - 0.1, high=constants.WORKSPACE_BOUNDS[1] + 0.1),
            instruction=spaces.Box(low=0, high=2147483647, shape=(constants.INSTRUCTION_LENGTH,), dtype=np.int32),
            rgb=spaces.Box(low=0, high=255, shape=(constants.IMAGE_HEIGHT, constants.IMAGE_WIDTH, 3), dtype=np.uint8)))

    # ------------------------------------------------------------------------------------------- helpers
    @staticmethod
    def encode_instruction(instruction: str) -> np.ndarray:
        raw = np.frombuffer(instruction.encode("utf-8"), dtype=np.uint8).astype(np.int32)[:constants.INSTRUCTION_LENGTH]
        out = np.zeros(constants.INSTRUCTION_LENGTH, np.int32)
        out[:len(raw)] = raw
        return out

    @staticmethod
    def decode_instruction(bytes_array) -> str:
        arr = np.asarray(bytes_array).astype(np.uint8)
        return bytes(arr[arr != 0].tolist()).decode("utf-8")

    def seed(self, seed=None):
        self._rng = np.random.RandomState(seed=seed)
        if self._reward_calculator is not None:
            self._reward_calculator.seed(self._rng)

    @property
    def succeeded(self) -> bool:
        if self._reward_calculator is None:
            return bool(self._board.succeeded[0])
        return bool(self._reward_calculator.reward(self._compute_state(request_task_update=False, render=False))[1])

    @property
    def instruction_str(self):
        return self._instruction_str

    @property
    def pybullet_client(self):
        """The reference hands out its Bullet client here (``:362``).  There is no physics engine behind this environment;
        the object returned answers the body / joint queries and resets of that interface from the board and the
        kinematic arm (``utils/board_client.py``), which is what ``ObjState.get_bullet_state / set_bullet_state`` need."""
        if getattr(self, "_board_client", None) is None:
            from language_table.environments.utils.board_client import BoardClient
            self._board_client = BoardClient(self)
        return self._board_client

    def calc_camera_params(self, image_size):
        """``(view matrix, projection matrix, position, look-at, up)`` of the table camera for ``image_size`` = (H, W), in
        the reference's column-major OpenGL convention (``:365-397``) — the math of :mod:`rt1b200.envs.camera`."""
        from rt1b200.envs import camera
        return camera.calc_camera_params(image_size, constants.CAMERA_POSE, constants.CAMERA_ORIENTATION, constants.CAMERA_INTRINSICS)

    def get_control_frequency(self):
        return self._control_frequency

    # ----------------------------------------------------------------------------------------- state / obs
    def _compute_state(self, request_task_update=True, render=True) -> Dict[str, Any]:
        """Everything the rewards, the oracles and the observation need, ``rgb`` last (the reference's ``_compute_state``
        renders too, ``:418-478``).  The frame comes from a one-entry cache keyed on the rendered quantities, so asking for
        the state of a scene that was just drawn costs no second rasterisation; ``render=False`` leaves ``rgb`` out."""
        b = self._board
        st = collections.OrderedDict(
            effector_translation=b.effector[0].cpu().numpy().astype(np.float32),
            effector_target_translation=b.effector_target[0].cpu().numpy().astype(np.float32),
            instruction=self._instruction)
        xy, yaw, mask = b.block_xy[0].cpu().numpy(), b.block_yaw[0].cpu().numpy(), b.block_mask[0].cpu().numpy()
        yaw = (yaw + np.pi) % (2.0 * np.pi) - np.pi              # reported like a quaternion's yaw: in [-pi, pi)
        for i, name in enumerate(self._all_blocks):
            st["block_%s_translation" % name] = xy[i].astype(np.float32)
            st["block_%s_orientation" % name] = np.array([yaw[i]], np.float32)
            st["block_%s_mask" % name] = np.array([1.0 if mask[i] else 0.0], np.float32)
        calc = self._reward_calculator
        if request_task_update and self._task_info is not None and hasattr(calc, "get_current_task_info"):
            # moving-target tasks (a side of another block, away from the others): the target follows the scene
            self._set_task_info(calc.get_current_task_info(st))
        self._update_obs_with_oracle_task_info(st)
        if render:
            st["rgb"] = self._render_camera()
        return st

    def _update_obs_with_oracle_task_info(self, st):
        """What the scripted oracle steers by (reference ``:509-540``): the block being pushed and the place it goes to, both
        relative to the effector TARGET; the place is the task's target point when it has one, else the target block."""
        e_target = st["effector_target_translation"]
        zero2, zero1 = np.zeros(2, np.float32), np.zeros(1, np.float32)
        if self._start_block is not None:
            st["effector_target_to_start_block_translation"] = (st["block_%s_translation" % self._start_block] - e_target).astype(np.float32)
            st["start_block_orientation"] = st["block_%s_orientation" % self._start_block]
        else:
            st["effector_target_to_start_block_translation"], st["start_block_orientation"] = zero2, zero1
        if self._oracle_target_translation is not None:
            st["effector_target_to_task_target_translation"] = (np.asarray(self._oracle_target_translation, np.float32) - e_target).astype(np.float32)
            st["task_target_orientation"] = zero1
        elif self._oracle_target_block is not None and "block_%s_translation" % self._oracle_target_block in st:
            st["effector_target_to_task_target_translation"] = (st["block_%s_translation" % self._oracle_target_block] - e_target).astype(np.float32)
            st["task_target_orientation"] = st["block_%s_orientation" % self._oracle_target_block]
        else:
            st["effector_target_to_task_target_translation"], st["task_target_orientation"] = zero2, zero1
        return st

    def _set_task_info(self, info):
        """Cache what defines the task (reference ``:742-816``): the pushed block, and where to — another block, a fixed point, or
        a point that moves with the scene — plus the instruction."""
        T = task_info
        target_block = location = direction = translation = None
        if isinstance(info, T.Block2BlockTaskInfo):
            start, target_block = info.block1, info.block2
        elif isinstance(info, T.Block2LocationTaskInfo):
            start, translation, location = info.block, info.target_translation, info.location
        elif isinstance(info, T.Block2LineTaskInfo):
            start, translation = info.block, info.target_translation
        elif isinstance(info, T.Block2RelativeLocationTaskInfo):
            start, translation, direction = info.block, info.target_translation, info.location
        elif isinstance(info, T.Block2BlockRelativeLocationTaskInfo):
            start, target_block, translation, direction = info.block, info.target_block, info.target_translation, info.direction
        elif isinstance(info, T.SeparateBlocksTaskInfo):
            start, translation = info.block, info.target_translation
        elif isinstance(info, T.Point2BlockTaskInfo):
            start = target_block = info.block_target
        elif isinstance(info, T.Block2PoleTaskInfo):
            start, target_block = info.block1, info.goal
        elif isinstance(info, T.PlayTaskInfo):
            start = None                                         # free-form instruction: nothing for an oracle to steer by
        else:
            raise ValueError("Unknown task info: %s" % (info,))
        if target_block is None and translation is None and not isinstance(info, T.PlayTaskInfo):
            raise ValueError("We are not providing the oracle with a target location to push a block to. Either the target block "
                             "or the target translation must be set by the reward calculator.")
        self._task_info = info
        self._start_block, self._oracle_target_block = start, target_block
        self._oracle_target_translation = None if translation is None else np.asarray(translation, np.float64)
        self._target_absolute_location, self._target_relative_location = location, direction
        if info.instruction != self._instruction_str:
            self._set_instruction(info.instruction)

    def compute_state(self, request_task_update=True):
        return self._compute_state(request_task_update)

    def _compute_observation(self, state=None):
        st = state or self.compute_state()
        return collections.OrderedDict((k, st[k]) for k in ("effector_translation", "effector_target_translation",
                                                              "instruction", "rgb"))

    def _render_camera(self) -> np.ndarray:
        b = self._board
        key = tuple(t[0].cpu().numpy().tobytes() for t in (b.effector, b.block_xy, b.block_yaw, b.block_mask))
        cached = getattr(self, "_frame_cache", None)
        if cached is None or cached[0] != key:
            cached = self._frame_cache = (key, b.render()[0].cpu().numpy())
        return cached[1].copy()

    def render(self, mode="rgb_array"):
        frame = self._render_camera()
        if not self._render_text_in_image:
            return frame
        return add_debug_info_to_image(frame, {"instruction": self._instruction_str} if self._instruction_str else {})

    # -------------------------------------------------------------------------------------- reset / step
    def reset(self, reset_poses=True):
        for _ in range(100):
            if reset_poses:
                subsets = blocks_module.get_all_block_subsets(self._block_mode, self._training) \
                    if self._block_mode == blocks_module.LanguageTableBlockVariants.N_CHOOSE_K else None
                self._board.gen.manual_seed(int(self._rng.randint(2 ** 31 - 1)))
                self._board.reset()
                if subsets is not None:
                    chosen = set(subsets[int(self._rng.randint(len(subsets)))])
                    self._board.block_mask[0] = torch.tensor([n in chosen for n in self._all_blocks])
                    # absent blocks are parked far off the board
                    self._board.block_xy[0][~self._board.block_mask[0]] = torch.tensor([5.0, 5.0])
                self._blocks_on_table = [n for n, m in zip(self._all_blocks, self._board.block_mask[0].tolist()) if m]
            if reset_poses and self._robot is not None:
                self._robot.reset_joints(constants.INITIAL_JOINT_POSITIONS)
                self._robot.reset_joints(self._robot.inverse_kinematics(self._effector_pose(self._board.effector[0])))
                # from here on the effector is wherever the joints put it (the IK residual is ~1e-6 m, but a saved
                # joint vector must reproduce the scene exactly)
                self._board.effector[0] = torch.as_tensor(np.asarray(self._robot.forward_kinematics().translation[:2], np.float32))
            state = self._compute_state(request_task_update=False, render=False)
            if self._reward_calculator is None:
                self._set_instruction("")
                break
            info = self._reward_calculator.reset(state, self._blocks_on_table)
            if info != task_info.FAILURE:
                self._set_task_info(info)
                break
        else:
            raise ValueError("could not sample a valid task for this board")
        return self._compute_observation()

    def _set_instruction(self, text: str):
        self._instruction_str = text
        self._instruction = self.encode_instruction(text)

    @staticmethod
    def _effector_pose(xy) -> Pose3d:
        xy = np.asarray(xy.cpu() if hasattr(xy, "cpu") else xy, np.float64)
        return Pose3d(rotation=constants.EFFECTOR_DOWN_ROTATION, translation=np.array([xy[0], xy[1], constants.EFFECTOR_HEIGHT]))

    @property
    def robot(self):
        return self._robot

    def _step_robot_and_sim(self, action) -> Optional[torch.Tensor]:
        """Command the arm to the new (workspace-clipped) target and run the control step's simulation sub-steps; returns the
        effector positions the board should sweep through, or ``None`` without an arm model."""
        if self._robot is None:
            return None
        lo, hi = constants.WORKSPACE_BOUNDS
        target = np.clip(self._board.effector_target[0].cpu().numpy().astype(np.float64) + np.asarray(action, np.float64), lo, hi)
        self._robot.set_target_effector_pose(self._effector_pose(target))
        n_wp = 4                                                # board contact resolution happens at 4 instants per step
        per = max(1, self._sim_steps_per_step // n_wp)
        path = []
        for _ in range(n_wp):
            self._robot.step_simulation(per, dt=1.0 / self._step_frequency)
            path.append(self._robot.forward_kinematics().translation[:2])
        return torch.as_tensor(np.asarray(path, np.float32))[None]

    def step(self, action):
        path = self._step_robot_and_sim(np.asarray(action, np.float32).reshape(2))
        action = torch.as_tensor(np.asarray(action, np.float32)).reshape(1, 2)
        self._board.step(action, effector_path=path)
        state = self._compute_state()
        reward, done = (0.0, False)
        if self._reward_calculator is not None:
            reward, done = self._reward_calculator.reward(state)
        return self._compute_observation(state), reward, done, {}

    # ------------------------------------------------------------------------------------ save / restore
    def get_board_state(self):
        """Snapshot of everything, the reward object and its task included (tensors, not serialisable): what a search or a
        what-if roll-out needs to come back to exactly this point."""
        b = self._board
        joints = None if self._robot is None else np.array(self._robot.get_joint_positions(), dtype=np.float64)
        return {"effector": b.effector.clone(), "effector_target": b.effector_target.clone(), "block_xy": b.block_xy.clone(),
                "block_yaw": b.block_yaw.clone(), "block_mask": b.block_mask.clone(), "instruction": self._instruction.copy(),
                "instruction_str": self._instruction_str, "blocks_on_table": list(self._blocks_on_table), "joints": joints,
                "reward": copy.deepcopy(self._reward_calculator), "task_info": copy.deepcopy(self._task_info)}

    def set_board_state(self, state):
        b = self._board
        b.effector, b.effector_target = state["effector"].clone(), state["effector_target"].clone()
        b.block_xy, b.block_yaw, b.block_mask = state["block_xy"].clone(), state["block_yaw"].clone(), state["block_mask"].clone()
        self._instruction, self._instruction_str = state["instruction"].copy(), state["instruction_str"]
        self._blocks_on_table = list(state["blocks_on_table"])
        if state.get("joints") is not None and self._robot is not None:
            self._robot.reset_joints(state["joints"])
            self._robot.set_target_effector_pose(self._effector_pose(b.effector_target[0]))
        if state.get("reward") is not None:
            self._reward_calculator = copy.deepcopy(state["reward"])
            self._reward_calculator.seed(self._rng)
        self._task_info = None
        self._start_block = self._oracle_target_block = self._oracle_target_translation = None
        self._target_absolute_location = self._target_relative_location = None
        if state.get("task_info") is not None:
            self._set_task_info(copy.deepcopy(state["task_info"]))

    def get_pybullet_state(self):
        """The scene as the reference saves it (``:234-294``): ``robots`` = [XarmState], ``robot_end_effectors`` /
        ``targets`` = [], ``objects`` = one ObjState per block, the oracle's task fields and the instruction, every string
        as its zero-padded byte code — plain data that ``utils_pybullet.write_pybullet_state`` can put in a file.  The
        reward object is not part of it (nor is it upstream); ``get_board_state`` keeps that too."""
        state = self.export_pybullet_state()
        for key, value in (("start_block", self._start_block), ("oracle_target_block", self._oracle_target_block),
                           ("target_absolute_location", self._target_absolute_location),
                           ("target_relative_location", self._target_relative_location)):
            if value is not None:
                state[key] = self.encode_instruction(str(value)).tolist()
        if self._oracle_target_translation is not None:
            state["oracle_target_translation"] = np.asarray(self._oracle_target_translation, np.float64).tolist()
        if self._instruction_str is not None:
            state["instruction_str"] = self.encode_instruction(self._instruction_str).tolist()
        return state

    def set_pybullet_state(self, state):
        """Restore a ``get_pybullet_state`` record (or one read back from a state file).  An instruction stored with a
        shorter code length is zero-padded to ``INSTRUCTION_LENGTH`` (reference ``:349-359``)."""
        if "block_xy" in state:                                  # a get_board_state snapshot
            return self.set_board_state(state)
        self.import_pybullet_state(state)
        for key, attr in (("start_block", "_start_block"), ("oracle_target_block", "_oracle_target_block"),
                          ("target_absolute_location", "_target_absolute_location"),
                          ("target_relative_location", "_target_relative_location")):
            if key in state:
                setattr(self, attr, self.decode_instruction(np.asarray(state[key])))
        if "oracle_target_translation" in state:
            self._oracle_target_translation = np.asarray(state["oracle_target_translation"], np.float64)
        if state.get("instruction_str"):
            self._instruction_str = self.decode_instruction(np.asarray(state["instruction_str"]))

    # ------------------------------------------------- the reference's state layout (for its gz-JSON state files)
    def export_pybullet_state(self) -> Dict[str, Any]:
        """The scene in the reference's ``get_pybullet_state`` layout (``:234-294``): ``robots`` = [XarmState] (joint angles,
        target effector pose), ``robot_end_effectors`` / ``targets`` = [], ``objects`` = one ObjState per block of the block
        mode (pose = ((x, y, z), quaternion); blocks that are off the table sit at the reference's parking height), plus the
        encoded instruction.  ``utils_pybullet.write_pybullet_state`` writes it to a file the reference can read."""
        from language_table.environments.utils import utils_pybullet as up
        client = self.pybullet_client
        objects = [up.ObjState.get_bullet_state(client, i) for i in range(len(self._all_blocks))]
        target = self._effector_pose(self._board.effector_target[0].tolist())
        if self._robot is not None:
            robot = up.XarmState.get_bullet_state(client, client.robot_id, target_effector_pose=target, goal_translation=None)
        else:
            robot = up.XarmState(obj_id=client.robot_id, base_pose=((0.0, 0.0, 0.0), (0.0, 0.0, 0.0, 1.0)),
                                 target_effector_pose=target, goal_translation=None)
        return {"robots": [robot], "robot_end_effectors": [], "targets": [], "objects": objects,
                "instruction": [int(v) for v in self._instruction.tolist()]}

    def import_pybullet_state(self, state: Dict[str, Any]) -> None:
        """Inverse of :meth:`export_pybullet_state` (block poses, arm joints, effector target, instruction)."""
        b = self._board
        client = self.pybullet_client
        for i, obj in enumerate(state["objects"][:len(self._all_blocks)]):
            obj.set_bullet_state(client, i)
        robot = state["robots"][0]
        if robot.target_effector_pose is not None:
            xy = torch.tensor(np.asarray(robot.target_effector_pose.translation[:2], np.float32))
            b.effector_target[0] = xy
            b.effector[0] = xy
        if self._robot is not None and len(robot.joint_state):
            if len(robot.joint_info) == len(robot.joint_state):
                robot.set_bullet_state(client, client.robot_id)
            else:                                               # a record without joint info (older files): positions only
                self._robot.reset_joints([js[0] for js in robot.joint_state])
            # the effector is where the restored joints put it (it lags the target within a control step)
            b.effector[0] = torch.as_tensor(np.asarray(self._robot.forward_kinematics().translation[:2], np.float32))
            if robot.target_effector_pose is not None:
                self._robot.set_target_effector_pose(robot.target_effector_pose)
        self._blocks_on_table = [n for n, m in zip(self._all_blocks, b.block_mask[0].tolist()) if m]
        if state.get("instruction") is not None:
            code = np.asarray(state["instruction"], np.int32)[:constants.INSTRUCTION_LENGTH]
            self._instruction = np.pad(code, (0, constants.INSTRUCTION_LENGTH - len(code)))
            self._instruction_str = self.decode_instruction(self._instruction)
