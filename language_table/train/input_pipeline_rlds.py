"""Input pipeline of the LAVA baseline (reference ``language_table/train/input_pipeline_rlds.py``).

The reference reads RLDS/TFDS episodes from GCS; offline the same step dictionaries come from this repository's per-episode
``.npy`` files (``rt1b200.data.episodes``, the format ``rlds_np_convert.py`` writes) or from the synthetic generator.
Kept behaviour: terminal steps dropped, episodes front-padded with copies of the first step so that every step ends a window
of ``sequence_length`` frames (``:105-176``), random crop of ``random_crop_factor`` of the area shared by the frames of a
window + bilinear resize to the target size (``:326-389``), photometric distortions (brightness / contrast / saturation,
``:391-458``), rgb scaled to [0, 1], actions normalised with the dataset statistics, a 512-d language embedding per step
under ``clip_embedding`` (the stored sentence embedding; the reference computes CLIP tokens/embeddings).
"""
import glob
import os
from typing import Dict, Iterator, List, Optional

import numpy as np
import torch
import torch.nn.functional as F


def _windows(episode: List[Dict], sequence_length: int):
    steps = [s for s in episode if not s.get("is_terminal", False)] or episode
    padded = [steps[0]] * (sequence_length - 1) + steps
    for end in range(sequence_length - 1, len(padded)):
        yield padded[end - sequence_length + 1:end + 1]


def random_resized_crop(frames: torch.Tensor, factor: float, out_hw, rng: np.random.Generator, train: bool) -> torch.Tensor:
    """frames uint8/float [t, h, w, 3] → float [t, H, W, 3]; one crop window per sequence (random in training, central in eval)."""
    t, h, w, _ = frames.shape
    ch, cw = max(1, int(round(h * np.sqrt(factor)))), max(1, int(round(w * np.sqrt(factor))))
    y0 = int(rng.integers(0, h - ch + 1)) if train else (h - ch) // 2
    x0 = int(rng.integers(0, w - cw + 1)) if train else (w - cw) // 2
    x = frames[:, y0:y0 + ch, x0:x0 + cw].permute(0, 3, 1, 2).float()
    x = F.interpolate(x, size=tuple(out_hw), mode="bilinear", align_corners=False)
    return x.permute(0, 2, 3, 1)


def photometric_distortions(frames: torch.Tensor, rng: np.random.Generator) -> torch.Tensor:
    """frames float [t, h, w, 3] in [0, 1]; the same random brightness / contrast / saturation for the whole window."""
    x = frames + float(rng.uniform(-0.125, 0.125))
    mean = x.mean(dim=(1, 2, 3), keepdim=True)
    x = (x - mean) * float(rng.uniform(0.5, 1.5)) + mean
    gray = x.mean(dim=-1, keepdim=True)
    x = (x - gray) * float(rng.uniform(0.5, 1.5)) + gray
    return x.clamp(0.0, 1.0)


class WindowDataset:
    """Iterable over (observation, action) windows of a list of episodes."""

    def __init__(self, episodes: List[List[Dict]], sequence_length: int, target_hw=(180, 320), random_crop_factor: Optional[float] = 0.95,
                 photometric: bool = True, train: bool = True, seed: int = 0):
        self.windows = [w for ep in episodes for w in _windows(ep, sequence_length)]
        self.hw, self.factor, self.photometric, self.train = tuple(target_hw), random_crop_factor, photometric, train
        self.rng = np.random.default_rng(seed)

    def __len__(self):
        return len(self.windows)

    def example(self, i: int) -> Dict:
        win = self.windows[i]
        rgb = torch.from_numpy(np.stack([np.asarray(s["rgb"]) for s in win]))
        rgb = random_resized_crop(rgb, self.factor or 1.0, self.hw, self.rng, self.train and self.factor is not None) / 255.0
        if self.train and self.photometric:
            rgb = photometric_distortions(rgb, self.rng)
        emb = torch.from_numpy(np.stack([np.asarray(s["instruction"], dtype=np.float32) for s in win]))
        eff = torch.from_numpy(np.stack([np.asarray(s.get("effector_translation", np.zeros(2)), dtype=np.float32) for s in win]))
        return {"observation": {"rgb": rgb, "clip_embedding": emb, "effector_translation": eff},
                "action": torch.from_numpy(np.asarray(win[-1]["action"], dtype=np.float32))}

    def batches(self, batch_size: int, shuffle: bool = True, drop_last: bool = True) -> Iterator[Dict]:
        order = self.rng.permutation(len(self)) if shuffle else np.arange(len(self))
        for i in range(0, len(order) - (batch_size - 1 if drop_last else 0), batch_size):
            ex = [self.example(int(j)) for j in order[i:i + batch_size]]
            yield {"observation": {k: torch.stack([e["observation"][k] for e in ex]) for k in ex[0]["observation"]},
                   "action": torch.stack([e["action"] for e in ex])}


def load_episodes(dataset_path: Optional[str], split: str = "train", synthetic: bool = False, num_synthetic: int = 4,
                  seed: int = 0) -> List[List[Dict]]:
    from rt1b200.data.episodes import load_episode, synthetic_episode
    if synthetic or not dataset_path:
        rng = np.random.default_rng(seed)
        return [synthetic_episode(rng, length=int(rng.integers(8, 16)), height=90, width=160) for _ in range(num_synthetic)]
    files = sorted(glob.glob(os.path.join(dataset_path, split, "episode_*.npy")))
    if not files:
        raise FileNotFoundError(f"no episode_*.npy under {os.path.join(dataset_path, split)}")
    return [load_episode(f) for f in files]


def _save_norm_info(path: str, info) -> None:
    os.makedirs(path, exist_ok=True)
    tmp = os.path.join(path, "norm_info.tmp.npy")
    np.save(tmp, np.array(info, dtype=object), allow_pickle=True)
    os.replace(tmp, os.path.join(path, "norm_info.npy"))                 # atomic: a waiting process never sees half a file


def _load_norm_info(path: str):
    f = os.path.join(path, "norm_info.npy")
    return np.load(f, allow_pickle=True).item() if os.path.exists(f) else None


def _wait_for_norm_info(path: str, timeout_s: float = 3600.0, poll_s: float = 0.2):
    import time
    t0 = time.time()
    while True:
        info = _load_norm_info(path)
        if info is not None:
            return info
        if time.time() - t0 > timeout_s:
            raise TimeoutError(f"process 0 did not write {path}/norm_info.npy within {timeout_s:.0f} s")
        time.sleep(poll_s)


def create_datasets(config, seed: int = 0, normalization_path: str = None, process_index: int = 0):
    """Returns ``(train WindowDataset, norm_info)`` like the reference's ``create_datasets`` (``:47-323``): statistics come from
    ``data_normalization_num_samples`` batches and actions are normalised with them.

    With ``normalization_path`` the statistics are handed off between processes the way the reference does it (``:191-233``,
    ``:517-529``): process 0 loads them from ``<normalization_path>/norm_info.npy`` or computes and writes them; every other
    process waits for the file and loads it, so all ranks (and a resumed run) normalise with the SAME numbers.  The data order
    of a process is seeded with ``seed + process_index`` (the reference folds ``jax.process_index()`` into its data rng,
    ``train.py:81``)."""
    from . import normalization
    eps = load_episodes(getattr(config, "dataset_path", None), "train", synthetic=getattr(config, "synthetic_data", False), seed=seed)
    ds = WindowDataset(eps, config.sequence_length, (config.data_target_height, config.data_target_width),
                       config.random_crop_factor if config.image_augment_crop else None, config.image_photometric_distortions,
                       train=True, seed=seed + process_index)
    if normalization_path is not None and process_index != 0:
        return ds, _wait_for_norm_info(normalization_path)
    info = _load_norm_info(normalization_path) if normalization_path is not None else None
    if info is None:
        stat_batches = ((b["observation"], b["action"]) for b in ds.batches(min(8, len(ds)), shuffle=False, drop_last=False))
        obs_stats, act_stats, amin, amax = normalization.compute_dataset_statistics(
            (({k: v.numpy() for k, v in o.items()}, a.numpy()) for o, a in stat_batches), config.data_normalization_num_samples,
            skip_keys=tuple(config.data_skip_normalize_keys) + ("clip_embedding",))
        info = {"observation_statistics": obs_stats, "action_statistics": act_stats, "action_min": amin, "action_max": amax}
        if normalization_path is not None:
            _save_norm_info(normalization_path, info)
    return ds, info
