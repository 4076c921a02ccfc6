"""Language-Table board, rewards, instruction grammar, oracles, wrappers and the closed-loop evaluator (CPU)."""
import os

import numpy as np
import pytest
import torch

from language_table.environments import blocks, constants, language_table
from language_table.environments.oracles import push_oracle_rrt_slowdown, rrt_star
from language_table.environments.rewards import (block1_to_corner, block2absolutelocation, block2block,
                                                 block2block_relative_location, block2relativelocation, instructions,
                                                 play, point2block, separate_blocks, task_info)
from language_table.environments.utils import pose3d, utils_pybullet, xarm_sim_robot


def test_instruction_counts_match_reference():
    # exact counts asserted by the reference's instructions_test.py:25-36
    V = blocks.LanguageTableBlockVariants
    assert len(instructions.generate_all_instructions(V.BLOCK_4)) == 12652
    assert len(instructions.generate_all_instructions(V.BLOCK_8)) == 30264
    assert len(instructions.generate_all_instructions(V.N_CHOOSE_K)) == 80368
    assert instructions.vocab_size(V.BLOCK_8) > 50


def test_block_sets():
    assert len(blocks.ALL_BLOCKS) == 16 and len(blocks.FIXED_8_COMBINATION) == 8
    assert len(blocks.TRAIN_COMBINATIONS) + len(blocks.TEST_COMBINATIONS) == len(blocks.ALL_COMBINATIONS)
    assert blocks.get_blocks_text_descriptions(blocks.LanguageTableBlockVariants.BLOCK_4)[0] == "red moon"


@pytest.mark.parametrize("mode", ["BLOCK_4", "BLOCK_8", "N_CHOOSE_K"])
def test_env_observation_and_step(mode):
    env = language_table.LanguageTable(blocks.LanguageTableBlockVariants(mode), block2block.BlockToBlockReward, seed=0)
    obs = env.reset()
    assert obs["rgb"].shape == (constants.IMAGE_HEIGHT, constants.IMAGE_WIDTH, 3) and obs["rgb"].dtype == np.uint8
    assert obs["instruction"].shape == (constants.INSTRUCTION_LENGTH,) and obs["instruction"].dtype == np.int32
    assert env.instruction_str == env.decode_instruction(obs["instruction"])
    for k in ("effector_translation", "effector_target_translation"):
        assert env.observation_space[k].contains(obs[k])
    before = obs["effector_target_translation"].copy()
    obs, reward, done, _ = env.step(np.array([0.02, -0.01], np.float32))
    np.testing.assert_allclose(obs["effector_target_translation"], before + [0.02, -0.01], atol=1e-6)
    assert reward == 0.0 and not done
    # the effector target never leaves the workspace
    for _ in range(60):
        obs, *_ = env.step(np.array([0.1, 0.1], np.float32))
    assert obs["effector_target_translation"][0] <= constants.X_MAX + 1e-6
    assert obs["effector_target_translation"][1] <= constants.Y_MAX + 1e-6


def test_state_save_restore_round_trip():
    env = language_table.LanguageTable(blocks.LanguageTableBlockVariants.BLOCK_8, block2block.BlockToBlockReward, seed=3)
    env.reset()
    saved = env.get_pybullet_state()
    img = env.render()
    for _ in range(5):
        env.step(np.array([0.03, 0.02], np.float32))
    assert np.abs(env.render().astype(float) - img).mean() > 0
    env.set_pybullet_state(saved)
    assert np.abs(env.render().astype(float) - img).mean() < 1e-6


@pytest.mark.parametrize("factory", [block2block.BlockToBlockReward, block2absolutelocation.BlockToAbsoluteLocationReward,
                                     block2relativelocation.BlockToRelativeLocationReward,
                                     block2block_relative_location.BlockToBlockRelativeLocationReward,
                                     block1_to_corner.Block1ToCornerLocationReward])
def test_oracle_solves_every_pushing_family(factory):
    env = language_table.LanguageTable(blocks.LanguageTableBlockVariants.BLOCK_8, factory, seed=1, render_text_in_image=False)
    env.reset()
    oracle = push_oracle_rrt_slowdown.ObstacleOrientedPushOracleBoard2dRRT(env, rng=np.random.RandomState(0))
    assert oracle.get_plan(env._compute_state())
    done = False
    for _ in range(200):
        env._board.step(torch.as_tensor(oracle.action(env._compute_state())).reshape(1, 2))
        reward, done = env._reward_calculator.reward(env._compute_state())
        if done:
            break
    assert done and reward == 100.0


def test_oracle_end_effector_planner_and_potential_fallback():
    """Reference push_oracle_rrt_slowdown.py:467-600: the effector's free-space route is planned around ALL blocks; when no
    route exists (effector boxed in by blocks) the potential field takes over and planning is retried."""
    O = push_oracle_rrt_slowdown
    env = language_table.LanguageTable(blocks.LanguageTableBlockVariants.BLOCK_8, block2block.BlockToBlockReward, seed=4,
                                       render_text_in_image=False)
    env.reset()
    b = env._board
    names = b.block_names
    start, target = names.index(env._task_info.block1), names.index(env._task_info.block2)
    # a wall of blocks between the effector and the pushed block: the direct line is blocked
    b.block_xy[0, start] = torch.tensor([0.45, 0.0])
    b.block_xy[0, target] = torch.tensor([0.45, 0.2])
    others = [i for i in range(len(names)) if i not in (start, target)]
    for k, i in enumerate(others):
        b.block_xy[0, i] = torch.tensor([0.33, -0.10 + 0.045 * k])
    b.effector[0] = b.effector_target[0] = torch.tensor([0.22, 0.0])
    oracle = O.ObstacleOrientedPushOracleBoard2dRRT(env, use_ee_planner=True, rng=np.random.RandomState(0))
    assert oracle.get_plan(env._compute_state()) and oracle.plan_success
    wall = b.block_xy[0, others].clone()
    done = False
    for _ in range(250):
        b.step(torch.as_tensor(oracle.action(env._compute_state())).reshape(1, 2))
        _, done = env._reward_calculator.reward(env._compute_state())
        if done:
            break
    assert done and not oracle.used_potential_fallback
    assert float((b.block_xy[0, others] - wall).norm(dim=-1).max()) < 0.02          # went around the wall, not through it
    # boxed in: four blocks hugging the effector leave RRT* no collision-free first step → potential-field fallback
    env.reset()
    b.effector[0] = b.effector_target[0] = torch.tensor([0.3, 0.0])
    ring = [(0.018, 0.0), (-0.018, 0.0), (0.0, 0.018), (0.0, -0.018)]
    free = [i for i in range(len(names)) if i not in (names.index(env._task_info.block1), names.index(env._task_info.block2))]
    for (dx, dy), i in zip(ring, free):
        b.block_xy[0, i] = torch.tensor([0.3 + dx, 0.0 + dy])
    oracle = O.ObstacleOrientedPushOracleBoard2dRRT(env, use_ee_planner=True, rng=np.random.RandomState(0))
    act = oracle.action(env._compute_state())
    assert oracle.used_potential_fallback and np.isfinite(act).all() and np.linalg.norm(act) <= 0.03 + 1e-6
    # helpers
    pts = O.filter_subgoals([np.array([0.0, 0.0]), np.array([0.005, 0]), np.array([0.03, 0]), np.array([0.032, 0]), np.array([0.06, 0])], 0.025)
    assert [round(float(p[0]), 3) for p in pts] == [0.0, 0.03, 0.06]
    assert O.slowdown_factor(0.01) == 0.2 and O.slowdown_factor(0.07) == 0.5 and O.slowdown_factor(0.5) == 1.0
    assert np.allclose(O.repulsive_push(np.array([0.0, 0.0]), np.array([0.1, 0.0])), 0)
    assert O.repulsive_push(np.array([0.0, 0.0]), np.array([0.04, 0.0]))[0] < 0 < O.attractive_pull(np.zeros(2), np.array([0.1, 0.0]))[0]


def test_point_separate_and_play_rewards():
    env = language_table.LanguageTable(blocks.LanguageTableBlockVariants.BLOCK_8, point2block.PointToBlockReward, seed=2)
    env.reset()
    tgt = env._compute_state()["block_%s_translation" % env._task_info.block_target]
    done = False
    for _ in range(100):
        st = env._compute_state()
        _, _, done, _ = env.step(np.clip(tgt - st["effector_target_translation"], -0.03, 0.03))
        if done:
            break
    assert done
    # separate: put two blocks next to each other so the task exists
    env = language_table.LanguageTable(blocks.LanguageTableBlockVariants.BLOCK_4, None, seed=0)
    env.reset()
    env._board.block_xy[0, 1] = env._board.block_xy[0, 0] + torch.tensor([0.045, 0.0])
    rew = separate_blocks.SeparateBlocksReward(100.0, np.random.RandomState(0), 0, blocks.LanguageTableBlockVariants.BLOCK_4)
    info = rew.reset(env._compute_state(), list(blocks.FIXED_4_COMBINATION))
    assert info == task_info.FAILURE or "from the" in info.instruction
    env8 = language_table.LanguageTable(blocks.LanguageTableBlockVariants.BLOCK_8, None, seed=0)
    env8.reset()
    pr = play.PlayReward(100.0, np.random.RandomState(0), 0, blocks.LanguageTableBlockVariants.BLOCK_8)
    info = pr.reset(env8._compute_state(), list(blocks.FIXED_8_COMBINATION))
    assert isinstance(info.instruction, str) and info.block1 != info.block2 and pr.reward({}) == (0.0, False)


def test_play_instruction_families():
    """Family sizes of the reference's long-horizon vocabulary (rewards/play.py): 100 training + 25 held-out 4-block
    instructions, eleven 8-block families sampled family-first."""
    train, test = play.get_100_4block_instructions(), play.get_100_4block_instructions(return_train=False)
    assert len(train) == 100 and len(test) == 25 and not set(train) & set(test)
    assert train == play.get_100_4block_instructions()                       # seeded split, no global random state touched
    assert [len(f) for f in play.four_block_families()] == [29, 288, 216, 72, 48]
    sizes = {f.__name__: len(f()) for f in play.EIGHT_BLOCK_FAMILIES if f is not play.blocks_in_order_outer_edge}
    assert sizes == {"get_sort_tasks": 1, "colors_in_locations": 72576, "group_color_pairs": 24, "get_colors_in_lines": 24,
                     "group_color_pairs_in_locations": 432, "get_line_tasks": 11, "get_surround_tasks": 8,
                     "all_blocks_in_location": 9, "k_blocks_in_location_i_rest_in_location_j": 504,
                     "get_shape_instructions": 22}
    rng = np.random.RandomState(0)
    draws = [play.get_random_8block_instruction(rng) for _ in range(400)]
    assert all(len(d) <= play.MAX_INSTRUCTION_LENGTH for d in draws)
    # family-first sampling: the single-instruction family shows up about 1 time in 11, not 1 in 114 000
    assert 15 <= draws.count("group the blocks by color") <= 70
    edge = play.sample_outer_edge(rng)
    assert edge.startswith("put the ") and all(b in edge for b in play.BLOCKS8)
    pr4 = play.PlayReward(100.0, np.random.RandomState(0), 0, blocks.LanguageTableBlockVariants.BLOCK_4)
    assert pr4._sample_instruction() in train


def test_rrt_star_routes_around_an_obstacle():
    planner = rrt_star.RRTStar([0.2, 0.0], [0.5, 0.0], [(0.35, 0.0, 0.06)], ((0.15, -0.3), (0.6, 0.3)),
                               rng=np.random.RandomState(0), max_iter=2048)
    path = planner.plan()
    assert path is not None and np.allclose(path[0], [0.2, 0.0]) and np.allclose(path[-1], [0.5, 0.0])
    d = np.linalg.norm(path[:, None] - np.array([[0.35, 0.0]]), axis=-1)
    assert d.min() > 0.055


def test_robot_and_state_utils(tmp_path):
    """xArm6 kinematics restated without PyBullet: the figures of the reference's xarm_sim_robot_test.py:50-78."""
    import math
    from scipy.spatial import transform
    from language_table.environments import constants
    robot = xarm_sim_robot.XArmSimRobot()
    robot.reset_joints([0, math.pi / 2, math.pi, 0, 0, 0])                  # pointing down the x axis
    x, y, _ = robot.forward_kinematics().translation
    assert abs(x - 0.714479) < 1e-3 and abs(y + 0.0006) < 1e-3
    robot.reset_joints([math.pi / 2, math.pi / 2, math.pi, 0, 0, 0])        # pointing down the y axis
    x, y, _ = robot.forward_kinematics().translation
    assert abs(x - 0.0006) < 1e-3 and abs(y - 0.714479) < 1e-3
    # inverse kinematics round trip (reference :64-78)
    robot = xarm_sim_robot.XArmSimRobot()
    initial = robot.forward_kinematics()
    target = pose3d.Pose3d(rotation=transform.Rotation.from_rotvec([0, math.pi / 2, 0]), translation=np.array([0.5, 0.0, 0.10]))
    robot.reset_joints(robot.inverse_kinematics(target))
    got = robot.forward_kinematics()
    assert not np.all(initial.vec7 == got.vec7)
    np.testing.assert_almost_equal(got.vec7, target.vec7, decimal=2)
    # the environment's initial joints are the tool-down pose at (0.3, -0.2, 0.145) that constants.py documents
    robot = xarm_sim_robot.XArmSimRobot(initial_joint_positions=constants.INITIAL_JOINT_POSITIONS)
    p0 = robot.forward_kinematics(link_frame=True)
    np.testing.assert_allclose(p0.translation, [0.3, -0.2, 0.145], atol=2e-3)
    np.testing.assert_allclose(np.abs(p0.rotation.as_rotvec()), [0, math.pi, 0], atol=2e-2)
    # position control: the joints servo towards the IK solution, so the effector lags its target, then converges
    pose = pose3d.Pose3d(rotation=constants.EFFECTOR_DOWN_ROTATION, translation=np.array([0.34, -0.17, constants.EFFECTOR_HEIGHT]))
    robot.set_target_effector_pose(pose)
    start = robot.forward_kinematics().translation.copy()
    robot.step_simulation(3)
    mid = robot.forward_kinematics().translation
    assert 1e-4 < np.linalg.norm(mid - start) < np.linalg.norm(pose.translation - start) - 1e-3
    robot.step_simulation(60)
    np.testing.assert_allclose(robot.forward_kinematics().translation, pose.translation, atol=5e-4)
    assert pose3d.Pose3d.deserialize(pose.serialize()) == pose
    st = utils_pybullet.XarmState(obj_id=1, base_pose=((0, 0, 0), (0, 0, 0, 1)), target_effector_pose=pose, goal_translation=np.zeros(3))
    path = str(tmp_path / "state.json.gz")
    utils_pybullet.write_state_file(path, [st.serialize()])
    assert utils_pybullet.read_state_file(path)[0]["obj_id"] == 1


def test_camera_matrices_and_view_rays():
    """View / projection matrices (reference language_table.py:365-397) and the ray helpers of utils_pybullet.py:115-196."""
    import torch
    from language_table.environments import constants as C
    from rt1b200.envs import camera
    from rt1b200.envs.scene import SceneBuilder
    viewm, projm, pos, lookat, up = camera.calc_camera_params((C.IMAGE_HEIGHT, C.IMAGE_WIDTH), C.CAMERA_POSE, C.CAMERA_ORIENTATION,
                                                              C.CAMERA_INTRINSICS)
    assert len(viewm) == 16 and len(projm) == 16
    pts = np.array([[0.3, 0.1, 0.0], [0.5, -0.2, 0.02], [0.2, 0.0, 0.04]])
    px = camera.project_points(pts, viewm, projm, C.IMAGE_WIDTH, C.IMAGE_HEIGHT)
    # the OpenGL matrices and the scene's pinhole projection are the same camera
    sb = SceneBuilder(["red_moon"], use_assets=False)
    np.testing.assert_allclose(px, sb.project(torch.tensor(pts, dtype=torch.float32))[:, :2].numpy(), atol=2e-3)
    # pixel → view ray → world ray → board plane recovers a point on the board
    hit = utils_pybullet.pixel_to_workspace(px[0], viewm, projm, C.IMAGE_WIDTH, C.IMAGE_HEIGHT)
    np.testing.assert_allclose(hit, pts[0], atol=1e-6)
    xyz, quat = utils_pybullet.decompose_view_matrix(viewm)
    np.testing.assert_allclose(xyz, C.CAMERA_POSE, atol=1e-9)
    assert abs(np.linalg.norm(quat) - 1) < 1e-9
    # an object at the look-at point sits on the camera's -z axis, one unit away
    v_xyz, _ = utils_pybullet.world_obj_to_view(lookat, [0, 0, 0, 1], viewm)
    np.testing.assert_allclose(v_xyz, [0, 0, -1], atol=1e-6)
    assert utils_pybullet.ray_to_plane_test([0, 0, 1], [1, 0, 0], [0, 0, 0], [0, 0, 1]) is None        # parallel
    assert utils_pybullet.ray_to_plane_test([0, 0, 1], [0, 0, 1], [0, 0, 0], [0, 0, 1]) is None        # behind the origin
    np.testing.assert_allclose(utils_pybullet.ray_to_plane_test([0, 0, 1], [0, 0, -1], [0, 0, 0], [0, 0, 1]), [0, 0, 0])


def test_env_effector_lags_target_within_a_control_step():
    env = language_table.LanguageTable(blocks.LanguageTableBlockVariants.BLOCK_4, block2block.BlockToBlockReward, seed=3)
    obs = env.reset()
    np.testing.assert_allclose(obs["effector_translation"], obs["effector_target_translation"], atol=1e-4)
    np.testing.assert_allclose(env.robot.forward_kinematics().translation[:2], obs["effector_translation"], atol=1e-4)
    obs, _, _, _ = env.step(np.array([0.03, -0.02], np.float32))
    lag = np.linalg.norm(obs["effector_target_translation"] - obs["effector_translation"])
    assert 1e-6 < lag < 5e-3                                            # behind the commanded target, but nearly there
    no_arm = language_table.LanguageTable(blocks.LanguageTableBlockVariants.BLOCK_4, block2block.BlockToBlockReward, seed=3,
                                          arm_kinematics=False)
    o2 = no_arm.reset()
    o2, _, _, _ = no_arm.step(np.array([0.03, -0.02], np.float32))
    np.testing.assert_allclose(o2["effector_translation"], o2["effector_target_translation"], atol=1e-7)


def test_wrappers_policy_and_entry_point(tmp_path):
    from language_table.eval import main_rt1, wrappers
    env = language_table.LanguageTable(blocks.LanguageTableBlockVariants.BLOCK_8, block2block.BlockToBlockReward, seed=0)
    env = wrappers.HistoryWrapper(wrappers.CentralCropImageWrapper(wrappers.UseTokenWrapper(env), 64, 96, 0.95), history_length=3)
    obs = env.reset()
    assert obs["rgb_sequence"].shape == (3, 64, 96, 3) and obs["instruction_tokenized_use"].shape == (3, 512)
    assert np.allclose(obs["rgb_sequence"][0], obs["rgb_sequence"][2])      # first observation is tiled
    assert abs(np.linalg.norm(obs["instruction_tokenized_use"][0]) - 1.0) < 1e-4
    res = main_rt1.main(["--workdir", str(tmp_path), "--episodes", "1", "--max_episode_steps", "2", "--height", "64", "--width", "96",
                         "--seq_len", "2", "--device", "cpu"])
    assert res["blocktoblock"]["episodes"] == 1
    # the other task families of the reference's table, driven by the oracle through the same loop; one video per episode
    res = main_rt1.main(["--workdir", str(tmp_path / "fam"), "--episodes", "1", "--max_episode_steps", "80", "--height", "64", "--width", "96",
                         "--seq_len", "2", "--device", "cpu", "--oracle", "--rewards", "block1tocorner,blocktoabsolutelocation",
                         "--block_mode", "BLOCK_4", "--seed", "1"])
    assert set(res) == {"block1tocorner", "blocktoabsolutelocation"} and res["block1tocorner"]["successes"] == 1
    assert sorted(os.listdir(tmp_path / "fam")) == ["block1tocorner_0.png", "blocktoabsolutelocation_0.png", "results.json"]
    with pytest.raises(ValueError):
        main_rt1.main(["--workdir", str(tmp_path / "bad"), "--device", "cpu", "--rewards", "nosuchtask"])


def test_batched_board_oracle_and_closed_loop_cpu():
    from rt1b200.envs.board import BatchedLanguageTable
    from rt1b200.eval.closed_loop import run_closed_loop
    env = BatchedLanguageTable(6, seed=1)
    for _ in range(80):
        env.step(env.scripted_action())
    assert int(env.succeeded.sum()) >= 5
    res = run_closed_loop(None, num_envs=3, episodes=1, max_episode_steps=80, device="cpu", use_oracle=True, image_hw=(64, 96)) \
        if False else None   # the policy-free path needs a model for dtype; covered in the GPU test


def test_compat_import_paths():
    import subprocess, sys, os
    repo = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = ("from pytorch_robotics_transformer.transformer_network import TransformerNetwork;"
            "from pytorch_robotics_transformer.tokenizers.action_tokenizer import RT1ActionTokenizer;"
            "from load_np_dataset import EmbodiedIntelligenceDataset; import distribute_train;"
            "assert distribute_train.RT1_Lightning is distribute_train.RT1Module; print('ok')")
    r = subprocess.run([sys.executable, "-c", code], cwd=repo, capture_output=True, text=True)
    assert r.returncode == 0 and "ok" in r.stdout, r.stderr[-2000:]


def test_render_writes_the_instruction_above_the_frame():
    """``render()`` = frame at 640x360 under a white band with the wrapped instruction (reference language_table.py:192-199,
    1000-1030); ``render_text_in_image=False`` returns the bare camera frame."""
    from language_table.environments import blocks, language_table
    from language_table.environments.rewards import block2block
    env = language_table.LanguageTable(block_mode=blocks.LanguageTableBlockVariants.BLOCK_8,
                                       reward_factory=block2block.BlockToBlockReward, seed=0)
    obs = env.reset()
    frame = env.render()
    assert frame.shape == (444, 640, 3) and frame.dtype == np.uint8
    band = frame[:84]
    assert (band == 255).mean() > 0.7 and (band < 128).any()           # white band with dark glyphs
    bare = language_table.LanguageTable(block_mode=blocks.LanguageTableBlockVariants.BLOCK_8,
                                        reward_factory=block2block.BlockToBlockReward, seed=0, render_text_in_image=False)
    bare.reset()
    assert bare.render().shape == obs["rgb"].shape
    empty = language_table.add_debug_info_to_image(obs["rgb"], {})
    assert empty.shape == (444, 640, 3) and (empty[:84] == 255).all()


def test_invisible_walls_keep_blocks_on_the_board():
    """``add_invisible_walls=True`` (reference language_table.py:933-990): a block pushed towards the edge stops at the wall."""
    import torch
    from language_table.environments import constants as C
    from rt1b200.envs.board import BatchedLanguageTable
    res = {}
    for walls in (False, True):
        b = BatchedLanguageTable(1, device="cpu", seed=0, add_invisible_walls=walls)
        b.reset()
        b.block_xy[0, 0] = torch.tensor([C.X_MAX - 0.08, 0.0])
        b.block_xy[0, 1:] = torch.tensor([C.X_MIN + 0.05, C.Y_MIN + 0.05])        # out of the way
        b.effector[0] = b.effector_target[0] = torch.tensor([C.X_MAX - 0.14, 0.0])
        for _ in range(12):
            b.step(torch.tensor([[0.02, 0.0]]))
        res[walls] = float(b.block_xy[0, 0, 0])
    assert res[True] <= C.X_MAX - 0.05 - C.BLOCK_RADIUS + 1e-6
    assert res[False] > res[True]


def test_reference_format_state_file_round_trip(tmp_path):
    """Scene → the reference's state layout (robots / objects as XarmState / ObjState) → gz-JSON file (version 2) → scene."""
    from language_table.environments import blocks, language_table
    from language_table.environments.rewards import block2block
    from language_table.environments.utils import utils_pybullet as up
    env = language_table.LanguageTable(block_mode=blocks.LanguageTableBlockVariants.BLOCK_4,
                                       reward_factory=block2block.BlockToBlockReward, seed=3)
    env.reset()
    for _ in range(3):
        env.step(np.array([0.01, -0.01], np.float32))
    st = env.export_pybullet_state()
    assert set(st) >= {"robots", "robot_end_effectors", "targets", "objects"} and isinstance(st["robots"][0], up.XarmState)
    f = str(tmp_path / "state.json.gz")
    up.write_pybullet_state(f, st, task="block2block", actions=[[0.01, -0.01]] * 3)
    data = up.read_pybullet_state(f)
    assert data["state_version"] == 2 and data["task"] == "block2block" and len(data["actions"]) == 3
    other = language_table.LanguageTable(block_mode=blocks.LanguageTableBlockVariants.BLOCK_4,
                                         reward_factory=block2block.BlockToBlockReward, seed=99)
    other.reset()
    other.import_pybullet_state(data["pybullet_state"])
    a, b = env._board, other._board
    assert np.allclose(a.block_xy.numpy(), b.block_xy.numpy(), atol=1e-6)
    assert np.allclose(np.cos(a.block_yaw.numpy() - b.block_yaw.numpy()), 1.0, atol=1e-6)      # same angle modulo 2 pi
    assert np.allclose(a.effector_target.numpy(), b.effector_target.numpy(), atol=1e-6)
    assert other.instruction_str == env.instruction_str
    assert np.allclose(other.robot.get_joint_positions(), env.robot.get_joint_positions())
    import gzip, json
    raw = json.loads(gzip.open(f, "rb").read().decode())
    assert raw["pybullet_state"]["objects"][0]["_serialized_obj_name"] == "ObjState"     # the reference's tagging


def test_rrt_entry_point_predicates_and_debug_plot(tmp_path):
    """The planner's public entry point and helpers (reference oracles/rrt_star.py:25-85, :275-332; plot.py:28)."""
    from language_table.environments.oracles import plot
    path, ok = rrt_star.get_shortest_path_no_collisions(
        [0.2, 0.0], [0.5, 0.0], (0.15, 0.6), (-0.3, 0.3), [(0.35, 0.0)], [0.05], delta=0.015, step_length=0.05,
        goal_sample_rate=0.1, search_radius=0.5, iter_max=1024, rng=np.random.RandomState(0))
    assert ok and np.allclose(path[0], [0.5, 0.0]) and np.allclose(path[-1], [0.2, 0.0])       # goal first
    for a, b in zip(path[:-1], path[1:]):
        assert not rrt_star.is_collision(a, b, 0.0, [(0.35, 0.0, 0.064)], [])
    # boxed in: no route, the straight line is handed back (or an error on request)
    ring = [(0.2 + 0.05 * np.cos(t), 0.05 * np.sin(t)) for t in np.linspace(0, 2 * np.pi, 12, endpoint=False)]
    args = dict(x_range=(0.15, 0.6), y_range=(-0.3, 0.3), obstacle_xy=ring, obstacle_widths=[0.02] * 12, delta=0.005,
                step_length=0.05, goal_sample_rate=0.1, search_radius=0.5, iter_max=64)
    path, ok = rrt_star.get_shortest_path_no_collisions([0.2, 0.0], [0.5, 0.0], **args)
    assert not ok and len(path) == 2 and np.allclose(path[0], [0.5, 0.0])
    with pytest.raises(ValueError):
        rrt_star.get_shortest_path_no_collisions([0.2, 0.0], [0.5, 0.0], raise_error_on_plan_failure=True, **args)
    assert rrt_star.is_collision([0.2, 0.0], [0.5, 0.0], 0.015, [(0.35, 0.0, 0.05)], [])
    assert not rrt_star.is_collision([0.2, 0.2], [0.5, 0.2], 0.015, [(0.35, 0.0, 0.05)], [])
    assert rrt_star.inside_boundary([0.15, 0.0], 0.01, [(0.15, -0.3, 0.01, 0.6)])
    assert rrt_star.is_intersect_circle([0, 0], [1, 0], [0.5, 0.05], 0.04, 0.02)
    assert not rrt_star.is_intersect_circle([0, 0], [1, 0], [1.5, 0.0], 0.04, 0.02)            # foot point beyond the end
    assert np.allclose(rrt_star.segment_distances([0, 0], [1, 0], [[0.5, 0.3], [2.0, 0.0]]), [0.3, 1.0])
    planner = rrt_star.RRTStar([0.2, 0.0], [0.5, 0.0], [(0.35, 0.0, 0.06)], ((0.15, -0.3), (0.6, 0.3)),
                               rng=np.random.RandomState(0))
    route = planner.plan()
    p = plot.PlotRRT([0.2, 0.0], [0.5, 0.0], [(0.15, -0.3, 0.45, 0.6)], [(0.35, 0.0, 0.06)], image_path=str(tmp_path / "rrt.png"))
    p.animation(planner, route, "rrt*")
    out = p.save_debug_image()
    assert open(out, "rb").read(8) == b"\x89PNG\r\n\x1a\n" and p.fig.shape == (400, 400, 3)
    assert (p.fig == np.array([220, 30, 30], np.uint8)).all(-1).any()                           # the path is drawn


def test_oriented_oracle_pushing_info():
    from language_table.environments.oracles.oriented_push_oracle import OrientedPushOracle, PushingInfo
    env = language_table.LanguageTable(blocks.LanguageTableBlockVariants.BLOCK_8, block2block.BlockToBlockReward, seed=0)
    env.reset()
    oracle = OrientedPushOracle(env)
    info = oracle.get_pushing_info()
    assert isinstance(info, PushingInfo) and len(info.obstacle_poses) == 7
    assert abs(np.linalg.norm(info.xy_dir_block_to_target) - 1) < 1e-6 and -np.pi <= info.theta_error <= np.pi
    # behind the block, on the line to the target: no orientation error
    state = dict(env._compute_state())
    state["effector_target_translation"] = info.xy_pre_block
    assert abs(oracle.get_pushing_info(state).theta_error) < 1e-6
    assert np.allclose(oracle.rotate(np.pi / 2, [1.0, 0.0]), [0.0, 1.0]) and abs(oracle.get_theta_from_vector([0, 2]) - np.pi / 2) < 1e-9


def test_board_client_and_bullet_state_records(tmp_path):
    """ObjState / XarmState.get_bullet_state, set_bullet_state (reference utils_pybullet.py:264-315, :341-351) against the
    Bullet-style client the environment hands out."""
    env = language_table.LanguageTable(blocks.LanguageTableBlockVariants.BLOCK_4, block2block.BlockToBlockReward, seed=0)
    env.reset()
    client = env.pybullet_client
    assert client is env.pybullet_client and client.getNumBodies() == len(env._all_blocks) + 1
    on = [i for i in range(len(env._all_blocks)) if client.getBasePositionAndOrientation(i)[0][2] == 0.0]
    assert len(on) == 4                                                               # the others are parked below the table
    rec = utils_pybullet.ObjState.get_bullet_state(client, on[0])
    assert rec.joint_info == () and len(rec.base_pose[0]) == 3 and len(rec.base_pose[1]) == 4
    arm = utils_pybullet.XarmState.get_bullet_state(client, client.robot_id, env._effector_pose([0.3, 0.0]), np.array([0.1, 0.2, 0.0]))
    assert len(arm.joint_info) == len(arm.joint_state) == 6 and arm.joint_info[0][1] == "joint1" and arm.goal_translation == (0.1, 0.2, 0.0)
    # move things, then restore from the records
    before_xy, before_q = env._board.block_xy[0, on[0]].clone(), env._robot.get_joint_positions().copy()
    client.resetBasePositionAndOrientation(on[0], (0.4, 0.1, 0.0), (0.0, 0.0, 0.0, 1.0))
    client.resetJointState(client.robot_id, 2, targetValue=0.5)
    assert abs(client.getJointState(client.robot_id, 2)[0] - 0.5) < 1e-9
    assert np.allclose(client.getBasePositionAndOrientation(on[0])[0], (0.4, 0.1, 0.0), atol=1e-6)
    rec.set_bullet_state(client, on[0])
    arm.set_bullet_state(client, client.robot_id)
    assert torch.allclose(env._board.block_xy[0, on[0]], before_xy) and np.allclose(env._robot.get_joint_positions(), before_q)
    # a record of another body does not fit: joint count / info are checked first
    with pytest.raises(ValueError):
        arm.set_bullet_state(client, on[0])
    bad = dataclasses_replace(arm, joint_info=tuple((i,) + ("other",) + tuple(info[2:]) for i, info in enumerate(arm.joint_info)))
    with pytest.raises(ValueError):
        bad.set_bullet_state(client, client.robot_id)
    with pytest.raises(ValueError):
        client.getBasePositionAndOrientation(99)
    # through the serialised form and back
    again = utils_pybullet.XarmState.deserialize(json_roundtrip(arm.serialize()))
    assert again.joint_info == arm.joint_info and again.goal_translation == arm.goal_translation
    assert np.allclose(again.target_effector_pose.translation, arm.target_effector_pose.translation)
    again.set_bullet_state(client, client.robot_id)


def dataclasses_replace(obj, **kw):
    import dataclasses
    return dataclasses.replace(obj, **kw)


def json_roundtrip(x):
    import json
    return json.loads(json.dumps(x))


def test_rrt_star_tree_invariants():
    """After a full search every stored cost equals the parent's cost plus the edge (rewiring pushes savings down the
    subtree), children mirror the parent links, every edge is collision free, and the steps compose as documented."""
    planner = rrt_star.RRTStar([0.2, -0.2], [0.55, 0.2], [(0.35, 0.0, 0.07), (0.45, 0.15, 0.04)], ((0.15, -0.3), (0.6, 0.3)),
                               rng=np.random.RandomState(3), max_iter=600, search_until_max_iter=True)
    path = planner.plan()
    assert planner.success and path is planner.path and np.allclose(path[-1], [0.55, 0.2])
    for i, (p, par) in enumerate(zip(planner.nodes, planner.parent)):
        if par < 0:
            assert i == 0 and planner.cost(i) == 0.0
            continue
        assert abs(planner.cost(i) - planner.get_new_cost(par, p)) < 1e-9
        assert i in planner.children[par] and planner._free(planner.nodes[par], p)
    assert sum(len(c) for c in planner.children) == len(planner.nodes) - 1
    length = float(np.linalg.norm(np.diff(path, axis=0), axis=1).sum())
    end = planner.search_goal_parent()
    assert abs(length - planner.get_new_cost(end, planner.goal)) < 1e-9
    assert length < 1.35 * np.linalg.norm(np.array([0.55, 0.2]) - np.array([0.2, -0.2]))       # RRT* straightens the route
    # the steps on their own
    step = planner.new_state(np.array([0.0, 0.0]), np.array([1.0, 0.0]))
    assert np.allclose(step, [planner.expand_dis, 0.0]) and planner.new_state(step, step) is None
    assert rrt_star.RRTStar.nearest_neighbor([[0, 0], [1, 1], [2, 2]], [0.9, 1.2]) == 1
    d, a = rrt_star.RRTStar.get_distance_and_angle([0, 0], [0, 2])
    assert abs(d - 2) < 1e-12 and abs(a - np.pi / 2) < 1e-12
    n = planner.node(end)
    assert isinstance(n, rrt_star.Node) and (n.parent is None) == (planner.parent[end] < 0)
    q = rrt_star.FIFOQueue()
    q.put(1), q.put(2)
    assert q.get() == 1 and not q.empty() and q.get() == 2 and q.empty()


def test_state_carries_the_oracle_task_fields():
    """effector_target_to_start_block_translation & co. (reference language_table.py:509-540), kept current for moving-target
    tasks through get_current_task_info (``:457-466``), and carried by the saved state (``:265-284``)."""
    env = language_table.LanguageTable(blocks.LanguageTableBlockVariants.BLOCK_8, block2block.BlockToBlockReward, seed=2)
    env.reset()
    st = env._compute_state(render=False)
    info = env._task_info
    assert env._start_block == info.block1 and env._oracle_target_block == info.block2 and env._oracle_target_translation is None
    tgt = st["effector_target_translation"]
    assert np.allclose(st["effector_target_to_start_block_translation"], st["block_%s_translation" % info.block1] - tgt)
    assert np.allclose(st["effector_target_to_task_target_translation"], st["block_%s_translation" % info.block2] - tgt)
    assert np.allclose(st["start_block_orientation"], st["block_%s_orientation" % info.block1])
    assert st["task_target_orientation"].shape == (1,) and list(st)[-1] != "rgb" and list(env._compute_state())[-1] == "rgb"

    env = language_table.LanguageTable(blocks.LanguageTableBlockVariants.BLOCK_8,
                                       block2block_relative_location.BlockToBlockRelativeLocationReward, seed=2)
    env.reset()
    info = env._task_info
    st = env._compute_state(render=False)
    assert env._target_relative_location == info.direction and env._oracle_target_block == info.target_block
    assert np.allclose(st["effector_target_to_task_target_translation"], np.asarray(info.target_translation) - st["effector_target_translation"], atol=1e-6)
    # the target is defined relative to another block: move that block and the oracle's target follows
    i = env._all_blocks.index(info.target_block)
    env._board.block_xy[0, i] += torch.tensor([0.02, -0.01])
    moved = env._compute_state(render=False)
    delta = moved["effector_target_to_task_target_translation"] - st["effector_target_to_task_target_translation"]
    assert np.allclose(delta, [0.02, -0.01], atol=1e-6)
    # saved and restored with the scene
    saved = env.get_pybullet_state()
    assert env.decode_instruction(np.asarray(saved["start_block"])) == info.block
    assert env.decode_instruction(np.asarray(saved["target_relative_location"])) == info.direction and "oracle_target_translation" in saved
    other = language_table.LanguageTable(blocks.LanguageTableBlockVariants.BLOCK_8, block2block.BlockToBlockReward, seed=5)
    other.reset()
    other.set_pybullet_state(saved)
    assert other._start_block == info.block and other._oracle_target_block == info.target_block
    back = other._compute_state(request_task_update=False, render=False)
    assert np.allclose(back["effector_target_to_task_target_translation"], moved["effector_target_to_task_target_translation"], atol=1e-5)
    with pytest.raises(ValueError):
        env._set_task_info(object())
