"""Episode files in the reference's on-disk format, and a synthetic generator for them.

Format (reference ``rlds_np_convert.py:20-33`` / ``load_np_dataset.py:91-127``):
``<root>/{train,val,test}/episode_<i>.npy`` is a pickled python list of per-step dicts with at least
``rgb uint8[H0,W0,3]``, ``instruction f32[512]`` (sentence embedding), ``action f32[2]``, ``is_terminal``,
``is_first``; other RLDS keys ride along unused.

There is no network on the build/bench boxes, so the Language-Table RLDS dataset and the Universal Sentence
Encoder are unavailable; :func:`write_synthetic_dataset` emits Language-Table-shaped episodes (frames rendered
by the synthetic board of :mod:`rt1b200.envs` when available, else procedural blocks) with random 512-d
instruction embeddings.  :func:`convert_rlds` is the hook where a real RLDS iterator plugs in.
"""
from __future__ import annotations

import os
from typing import Callable, Dict, Iterable, List, Optional, Sequence

import numpy as np

STEP_KEYS = ("rgb", "instruction", "action", "is_terminal", "is_first", "is_last", "reward",
             "effector_translation", "effector_target_translation")


def make_step(rgb, instruction, action, is_first=False, is_last=False, is_terminal=False, reward=0.0,
              effector_translation=None, effector_target_translation=None) -> Dict:
    return {
        "rgb": np.asarray(rgb, dtype=np.uint8),
        "instruction": np.asarray(instruction, dtype=np.float32),
        "action": np.asarray(action, dtype=np.float32),
        "is_first": bool(is_first), "is_last": bool(is_last), "is_terminal": bool(is_terminal),
        "reward": np.float32(reward),
        "effector_translation": np.zeros(2, np.float32) if effector_translation is None
        else np.asarray(effector_translation, np.float32),
        "effector_target_translation": np.zeros(2, np.float32) if effector_target_translation is None
        else np.asarray(effector_target_translation, np.float32),
    }


def save_episode(path: str, steps: Sequence[Dict]) -> None:
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    arr = np.empty(len(steps), dtype=object)
    for i, s in enumerate(steps):
        arr[i] = s
    np.save(path, arr, allow_pickle=True)


def load_episode(path: str) -> List[Dict]:
    return list(np.load(path, allow_pickle=True))


def synthetic_episode(rng: np.random.Generator, length: int, height: int = 180, width: int = 320,
                      n_blocks: int = 8) -> List[Dict]:
    """A pushing episode on a procedural board: coloured squares, a round effector moving towards one of them."""
    instr = rng.standard_normal(512).astype(np.float32)
    instr /= max(1e-6, float(np.linalg.norm(instr)))
    blocks = rng.uniform([0.15, 0.15], [0.85, 0.85], size=(n_blocks, 2))
    colors = rng.integers(40, 255, size=(n_blocks, 3), dtype=np.uint8)
    eff = rng.uniform([0.2, 0.2], [0.8, 0.8])
    target = blocks[rng.integers(n_blocks)]
    yy, xx = np.mgrid[0:height, 0:width]
    steps = []
    for t in range(length):
        delta = np.clip((target - eff) * 0.25 + rng.normal(0, 0.004, 2), -0.1, 0.1)
        frame = np.full((height, width, 3), (196, 176, 150), np.uint8)
        for bpos, col in zip(blocks, colors):
            cy, cx = int(bpos[1] * height), int(bpos[0] * width)
            frame[max(0, cy - 7):cy + 7, max(0, cx - 7):cx + 7] = col
        ey, ex = eff[1] * height, eff[0] * width
        frame[(yy - ey) ** 2 + (xx - ex) ** 2 < 36] = (30, 30, 30)
        last = t == length - 1
        steps.append(make_step(frame, instr, delta.astype(np.float32), is_first=(t == 0), is_last=last,
                               is_terminal=last, effector_translation=eff, effector_target_translation=eff + delta))
        eff = np.clip(eff + delta, 0.0, 1.0)
    return steps


def write_synthetic_dataset(root: str, n_train: int = 8, n_val: int = 2, n_test: int = 2, seed: int = 0,
                            min_len: int = 8, max_len: int = 20, height: int = 180, width: int = 320) -> str:
    rng = np.random.default_rng(seed)
    for split, n in (("train", n_train), ("val", n_val), ("test", n_test)):
        for i in range(n):
            ep = synthetic_episode(rng, int(rng.integers(min_len, max_len + 1)), height, width)
            save_episode(os.path.join(root, split, f"episode_{i}.npy"), ep)
    return root


def convert_rlds(episodes: Iterable[Iterable[Dict]], out_root: str, embed: Callable[[str], np.ndarray],
                 splits=(("train", 7800), ("val", 100), ("test", 100)), decode=None) -> None:
    """RLDS → npy conversion (reference ``rlds_np_convert.py``: first 8 000 episodes split 7 800/100/100, the
    byte-encoded instruction replaced by its sentence embedding).  ``episodes`` yields episodes of RLDS step
    dicts (``observation`` with ``rgb``/``instruction`` bytes, ``action``, ``is_first`` …); ``embed`` maps the
    decoded instruction string to a 512-d vector (Universal Sentence Encoder upstream)."""
    decode = decode or (lambda b: bytes(int(x) for x in np.asarray(b).tolist() if x != 0).decode("utf-8"))
    it = iter(episodes)
    for split, count in splits:
        for i in range(count):
            try:
                ep = next(it)
            except StopIteration:
                return
            steps, cache = [], {}
            for st in ep:
                obs = st["observation"]
                text = decode(obs["instruction"])
                if text not in cache:
                    cache[text] = np.asarray(embed(text), np.float32)
                steps.append(make_step(obs["rgb"], cache[text], st["action"], st.get("is_first", False),
                                       st.get("is_last", False), st.get("is_terminal", False),
                                       st.get("reward", 0.0), obs.get("effector_translation"),
                                       obs.get("effector_target_translation")))
            save_episode(os.path.join(out_root, split, f"episode_{i}.npy"), steps)
