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


def _to_numpy(v):
    """RLDS/TFDS hands out eager tensors (``.numpy()``); plain numpy / python values pass through."""
    if hasattr(v, "numpy") and callable(v.numpy):
        v = v.numpy()
    return v


def decode_instruction(inst) -> str:
    """The Language-Table datasets store the instruction as zero-padded utf-8 byte values (reference ``rlds_np_convert.py:9-11``)."""
    if isinstance(inst, (bytes, str)):
        return inst.decode("utf-8") if isinstance(inst, bytes) else inst
    arr = np.asarray(inst)
    return bytes(int(x) for x in arr[arr != 0].tolist()).decode("utf-8")


def flatten_rlds_episode(raw_episode) -> List[Dict]:
    """One RLDS episode (``{"steps": iterable of {"observation": {...}, "action": ..., "is_first": ..., ...}}``, or the
    bare iterable of steps) → list of flat step dicts: every step-level key except ``observation`` plus every observation key,
    values as numpy — exactly what the reference's ``create_episode`` stores (``rlds_np_convert.py:13-26``)."""
    steps = raw_episode["steps"] if isinstance(raw_episode, dict) and "steps" in raw_episode else raw_episode
    out = []
    for step in steps:
        flat = {k: _to_numpy(v) for k, v in step.items() if k != "observation"}
        for k, v in step.get("observation", {}).items():
            flat[k] = _to_numpy(v)
        out.append(flat)
    return out


def convert_rlds(episodes: Iterable, out_root: str, embed: Callable[[str], np.ndarray],
                 splits=(("train", 7800), ("val", 100), ("test", 100)), raw_root: Optional[str] = None, decode=None) -> Dict[str, int]:
    """RLDS → npy conversion (reference ``rlds_np_convert.py``): the first ``sum(counts)`` episodes are split
    7 800 / 100 / 100; each becomes ``<out_root>/<split>/episode_<i>.npy`` with the byte-encoded instruction replaced by its
    512-d sentence embedding (Universal Sentence Encoder upstream — ``embed`` here), all other step and observation keys kept.
    With ``raw_root`` the un-embedded episodes are written too, as the reference does (its ``inst_*`` tree).  ``episodes`` is
    any iterable of RLDS episodes (TFDS dataset, or numpy exports with the same nesting).  Returns episodes written per split."""
    decode = decode or decode_instruction
    it = iter(episodes)
    written: Dict[str, int] = {}
    for split, count in splits:
        written[split] = 0
        for i in range(count):
            try:
                raw = next(it)
            except StopIteration:
                return written
            steps = flatten_rlds_episode(raw)
            if raw_root is not None:
                save_episode(os.path.join(raw_root, split, f"episode_{i}.npy"), steps)
            cache: Dict[str, np.ndarray] = {}
            embedded = []
            for st in steps:
                text = decode(st["instruction"])
                if text not in cache:
                    cache[text] = np.asarray(embed(text), np.float32).reshape(-1)
                st = dict(st)
                st["instruction"] = cache[text]
                embedded.append(st)
            save_episode(os.path.join(out_root, split, f"episode_{i}.npy"), embedded)
            written[split] += 1
    return written


def iter_exported_episodes(source_dir: str):
    """Episodes exported to disk with the RLDS nesting: ``<source_dir>/*.npy`` (pickled list of step dicts, or a dict with a
    ``steps`` entry), in name order — the offline stand-in for ``tfds.builder_from_directory(...).as_dataset()``."""
    for name in sorted(os.listdir(source_dir), key=lambda n: (len(n), n)):
        if name.endswith(".npy"):
            obj = np.load(os.path.join(source_dir, name), allow_pickle=True)
            yield obj.item() if obj.dtype == object and obj.shape == () else list(obj)
