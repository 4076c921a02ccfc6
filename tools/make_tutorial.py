"""Writes ``language_table/examples/language_table_tutorial.ipynb`` and, unless ``--no-run``, executes every code cell
in one namespace first, so a notebook that reaches the tree has run end to end on this machine (CPU is enough).

The reference ships a Colab tutorial for the simulator and the RLDS datasets
(``language_table/examples/language_table_tutorial.ipynb``: create an environment, look at ``obs['rgb']`` and the
instruction, step it, stream episodes from GCS, show them as videos).  This notebook covers the same ground with what
exists offline here: the PyBullet-free board, the scripted oracle, the APNG writer instead of mediapy, and the
``.npy`` episode format of ``rlds_np_convert.py`` instead of a GCS bucket — plus the step the reference leaves to the
reader, running the RT-1 policy in the loop.

    python tools/make_tutorial.py [--no-run]
"""
import argparse
import json
import os
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)

CELLS = []


def md(text):
    CELLS.append(("markdown", text.strip("\n")))


def code(text):
    CELLS.append(("code", text.strip("\n")))


md("""
# Language-Table with rt1b200: simulator, oracle, datasets, policy

Everything below runs offline and on a CPU; with a B200 the same calls run the CUDA rasteriser and the native policy
kernels (`device="cuda"`).  Run the notebook from the repository root, or put the root on `sys.path` first.
""")

code("""
import os, sys, tempfile
import numpy as np

if not os.path.isdir("language_table"):          # started from language_table/examples
    sys.path.insert(0, os.path.abspath(os.path.join("..", "..")))

from language_table.environments import blocks, language_table
from language_table.environments.rewards import block2block
from rt1b200.utils.video import write_apng

OUT = tempfile.mkdtemp(prefix="lt_tutorial_")


def show(frames, name, fps=5):
    \"\"\"One frame -> PNG, several -> animated PNG; displayed inline under Jupyter, otherwise only written to OUT.\"\"\"
    frames = np.asarray(frames)
    path = write_apng(os.path.join(OUT, name + ".png"), frames[None] if frames.ndim == 3 else frames, fps=fps)
    try:
        from IPython.display import Image, display
        display(Image(filename=path))
    except ImportError:
        print("wrote", path)
    return path
""")

md("""
## The environment

`LanguageTable` has the interface of the upstream class (`reset / step / render / succeeded / compute_state`).

**Block modes** — `BLOCK_8`: red moon, red pentagon, blue moon, blue cube, green cube, green star, yellow star, yellow
pentagon; `BLOCK_4`: red moon, blue cube, green star, yellow pentagon; `N_CHOOSE_K`: a random subset of the 12 possible
blocks.

**Rewards** (`language_table.environments.rewards`)

| class | task |
|---|---|
| `block2block.BlockToBlockReward` | push one block to another |
| `block2absolutelocation.BlockToAbsoluteLocationReward` | push a block to a named region of the board |
| `block2relativelocation.BlockToRelativeLocationReward` | push a block in a direction relative to where it is |
| `block2block_relative_location.BlockToBlockRelativeLocationReward` | push a block to a side of another block |
| `separate_blocks.SeparateBlocksReward` | push blocks apart |
| `point2block.PointToBlockReward` | move the effector to a block |
| `block1_to_corner.Block1ToCornerLocationReward` | single block to a corner of the board |
| `play.PlayReward` | samples one of the task families above per episode |
""")

code("""
def decode_inst(inst):
    \"\"\"The instruction is a zero-padded array of utf-8 bytes.\"\"\"
    inst = np.asarray(inst)
    return bytes(inst[inst != 0].astype(np.uint8).tolist()).decode("utf-8")


env = language_table.LanguageTable(block_mode=blocks.LanguageTableBlockVariants.BLOCK_8,
                                   reward_factory=block2block.BlockToBlockReward, seed=0)
obs = env.reset()
print({k: (v.shape, v.dtype) for k, v in obs.items()})
print(decode_inst(obs["instruction"]))
show(obs["rgb"], "first_observation")
show(env.render(), "first_observation_with_text")
""")

md("""
`env.render()` returns the camera image scaled to 640x360 under a white band with the instruction written into it, like the
upstream class (pass `render_text_in_image=False` for the bare frame).  An action is a 2-D effector displacement in metres, clipped
by the policy wrappers to ±0.03 per 10 Hz control step.  The arm model behind it (xArm6 forward / inverse kinematics and
a joint servo) makes the effector lag its target within a step, as in the upstream simulation.
""")

code("""
frames = []
for _ in range(8):
    obs, reward, done, info = env.step(np.array([0.02, 0.0], np.float32))
    frames.append(obs["rgb"])
print("effector", obs["effector_translation"], "target", obs["effector_target_translation"])
show(frames, "push_forward")
""")

md("""
## The scripted oracle

`ObstacleOrientedPushOracleBoard2dRRT` plans a block route with RRT*, an effector route to the pre-push pose (with a
potential-field fallback), and slows down near the goal.  It is what produced the `*_oracle_sim` datasets upstream, and
it is how episodes are generated here when no dataset is on disk.
""")

code("""
from language_table.environments.oracles import push_oracle_rrt_slowdown

oracle = push_oracle_rrt_slowdown.ObstacleOrientedPushOracleBoard2dRRT(env, rng=np.random.RandomState(0))
for _ in range(20):                               # re-sample the board until the oracle finds a plan
    obs = env.reset()
    oracle.reset()
    if oracle.get_plan(env._compute_state()):
        break
instruction = decode_inst(obs["instruction"])
episode, frames = [], [obs["rgb"]]
for t in range(60):
    action = np.asarray(oracle.action(env._compute_state()), np.float32)
    episode.append((obs, action))
    obs, reward, done, _ = env.step(action)
    frames.append(obs["rgb"])
    if done:
        break
print(instruction, "| steps", len(episode), "| succeeded", env.succeeded)
show(frames[::3], "oracle_episode")
""")

md("""
## Saving and restoring the scene

`env.get_pybullet_state()` returns the scene in the upstream layout — an `XarmState` for the arm (joint vector, controller set
point), one `ObjState` per block, the instruction — read through `env.pybullet_client`, a Bullet-style facade over the board.
`utils_pybullet.write_pybullet_state` puts it in a gz-JSON state file (format version 2, readable by the upstream tools);
`set_pybullet_state` brings an environment back to it, rendering included.
""")

code("""
from language_table.environments.utils import utils_pybullet

saved = env.get_pybullet_state()
before = env._compute_state()
state_file = os.path.join(OUT, "scene.json.gz")
utils_pybullet.write_pybullet_state(state_file, saved, task="block2block", actions=[a.tolist() for _, a in episode[:3]])
for _ in range(5):
    env.step(np.array([-0.03, 0.01], np.float32))
env.set_pybullet_state(utils_pybullet.read_pybullet_state(state_file)["pybullet_state"])
after = env._compute_state()
same = all(np.allclose(np.cos(before[k] - after[k]), 1.0, atol=1e-6) if k.endswith("_orientation")      # angles: modulo 2 pi
           else np.allclose(before[k], after[k], atol=1e-3) for k in before if k != "rgb")
print("restored:", same, "| mean pixel difference", float(np.abs(before["rgb"].astype(float) - after["rgb"]).mean()))
block_id = 0
print("block", block_id, "pose via the client:", env.pybullet_client.getBasePositionAndOrientation(block_id))
""")

md("""
## Episodes on disk

Training reads one `.npy` file per episode (`<root>/{train,val,test}/episode_<n>.npy`, an object array of step dicts with
`rgb`, `instruction` (a 512-d sentence embedding), `action`, `is_first / is_last / is_terminal`, …) — the format
`rlds_np_convert.py` writes from an RLDS / TFDS source.  TFDS and the GCS buckets of the upstream tutorial are not
reachable offline; the converter takes any iterable of RLDS-shaped episodes, so the oracle episode above goes through the
same code path.  (For a whole dataset: `python rlds_np_convert.py --out data --oracle 200 --train 180 --val 10 --test 10`
collects successful oracle episodes with `language_table.examples.collect_oracle_episodes.oracle_episodes` and converts them.)
""")

code("""
from rt1b200.data.episodes import convert_rlds, load_episode

rlds_like = [{"steps": [{"observation": {"rgb": o["rgb"], "instruction": o["instruction"],
                                         "effector_translation": o["effector_translation"],
                                         "effector_target_translation": o["effector_target_translation"]},
                         "action": a, "reward": np.float32(0.0), "is_first": i == 0, "is_last": i == len(episode) - 1,
                         "is_terminal": i == len(episode) - 1} for i, (o, a) in enumerate(episode)]}]


def embed(text):                                  # stands in for the Universal Sentence Encoder (512-d)
    rng = np.random.default_rng(abs(hash(text)) % (2 ** 32))
    return rng.standard_normal(512).astype(np.float32)


root = os.path.join(OUT, "np_dataset")
convert_rlds(rlds_like, root, embed, splits=(("train", 1),))
steps = load_episode(os.path.join(root, "train", "episode_0.npy"))
print(len(steps), "steps;", {k: np.asarray(v).shape for k, v in steps[0].items()})
""")

md("""
`load_np_dataset.EmbodiedIntelligenceDataset` (the class `distribute_train.py` uses) turns those files into training
windows of `seq_len` frames:
""")

code("""
from load_np_dataset import DecodeAndRandomResizedCrop, EmbodiedIntelligenceDataset

ds = EmbodiedIntelligenceDataset(os.path.join(root, "train"), [0], window_length=6,
                                 transform=DecodeAndRandomResizedCrop(0.95, (112, 64)))
sample = ds[len(ds) - 1]
print(len(ds), "windows;", {k: tuple(v.shape) for k, v in sample["train_observation"].items()},
      {k: tuple(v.shape) for k, v in sample["action_label"].items()})
""")

md("""
## RT-1 in the loop

`BCJaxPyPolicyRT1` adapts a `TransformerNetwork` to the environment (history of 6 frames, centre crop 0.95, resize,
action clipping).  A randomly initialised, shrunken model is used here so the cell runs in seconds on a CPU; with a
checkpoint from `distribute_train.py`, `python -m language_table.eval.main_rt1 --workdir out --ckpt <file>` does the
same for the full model and writes one video per episode.  On a GPU, `--num_envs 64` steps 64 boards per policy call.
""")

code("""
import torch
from language_table.eval import wrappers
from language_table.train.policy import BCJaxPyPolicyRT1
from rt1b200.models.factory import RT1Config, build_rt1

H, W, T = 64, 112, 6
model = build_rt1(RT1Config.language_table(height=H, width=W, seq_len=T, num_layers=1), eval_center_crop=True).eval()
wrapped = wrappers.HistoryWrapper(wrappers.CentralCropImageWrapper(wrappers.UseTokenWrapper(env), H, W, 0.95),
                                  history_length=T, tile_first_step_obs=True)
policy = BCJaxPyPolicyRT1(model, device="cpu")
policy.reset(1)
obs = wrapped.reset()
frames = [env.render()]
with torch.no_grad():
    for t in range(4):
        obs, reward, done, _ = wrapped.step(policy.action(obs))
        frames.append(env.render())
print("ran", len(frames) - 1, "policy steps; succeeded:", env.succeeded)
show(frames, "policy_rollout")
""")


def notebook():
    cells = []
    for kind, text in CELLS:
        lines = [l + "\n" for l in text.split("\n")]
        lines[-1] = lines[-1][:-1]
        cell = {"cell_type": kind, "metadata": {}, "source": lines}
        if kind == "code":
            cell.update({"execution_count": None, "outputs": []})
        cells.append(cell)
    return {"cells": cells, "metadata": {"kernelspec": {"display_name": "Python 3", "language": "python", "name": "python3"},
                                         "language_info": {"name": "python"}}, "nbformat": 4, "nbformat_minor": 5}


def run_cells():
    ns = {"__name__": "__tutorial__"}
    cwd = os.getcwd()
    os.chdir(REPO)
    try:
        for i, (kind, text) in enumerate(CELLS):
            if kind == "code":
                exec(compile(text, f"<cell {i}>", "exec"), ns)
    finally:
        os.chdir(cwd)
    return ns


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--no-run", action="store_true")
    a = ap.parse_args()
    if not a.no_run:
        run_cells()
    out = os.path.join(REPO, "language_table", "examples", "language_table_tutorial.ipynb")
    with open(out, "w") as f:
        json.dump(notebook(), f, indent=1)
    print("wrote", out)
