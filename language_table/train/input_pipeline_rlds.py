"""Input pipeline of the LAVA baseline (reference ``language_table/train/input_pipeline_rlds.py``).

The reference reads RLDS/TFDS episodes from GCS; offline the same step dictionaries come from this repository's per-episode
``.npy`` files (``rt1b200.data.episodes``, the format ``rlds_np_convert.py`` writes) or from the synthetic generator.
Kept behaviour: terminal steps dropped, episodes front-padded with copies of the first step so that every step ends a window
of ``sequence_length`` frames (``:105-176``), one random crop window of ``random_crop_factor`` of each side shared by the frames of a
sequence + bilinear resize to the target size (``:326-389``), photometric distortions (brightness / saturation / hue /
contrast, one level per sequence, ``:391-458``), rgb scaled to [0, 1], actions normalised with the dataset statistics, a 512-d language embedding per step
under ``clip_embedding`` (the stored sentence embedding; the reference computes CLIP tokens/embeddings).
"""
import dataclasses
import glob
import os
from typing import Dict, Iterator, List, Optional, Tuple

import numpy as np
import torch
import torch.nn.functional as F


def _windows(episode: List[Dict], sequence_length: int):
    steps = [s for s in episode if not s.get("is_terminal", False)] or episode
    padded = [steps[0]] * (sequence_length - 1) + steps
    for end in range(sequence_length - 1, len(padded)):
        yield padded[end - sequence_length + 1:end + 1]


def _decode_frames(frames) -> torch.Tensor:
    """uint8 array / tensor [t, h, w, 3] as it is; a sequence of encoded images (JPEG / PNG bytes, the RLDS storage
    format) through Pillow."""
    if torch.is_tensor(frames):
        return frames
    if isinstance(frames, np.ndarray) and frames.dtype != object and frames.dtype.kind not in "SU":
        return torch.from_numpy(frames)
    import io
    from PIL import Image
    return torch.from_numpy(np.stack([np.asarray(Image.open(io.BytesIO(bytes(f))).convert("RGB")) for f in frames]))


@dataclasses.dataclass(frozen=True)
class DecodeAndRandomResizedCrop:
    """Decode, cut one window of ``random_crop_factor`` × height by ``random_crop_factor`` × width (whole pixels, the same
    window for every frame of the sequence, its corner uniform over the offsets that keep it inside), scale to [0, 1],
    bilinear-resize to ``resize_size`` (reference ``input_pipeline_rlds.py:326-387``).  ``None`` keeps the whole frame.
    ``train=False`` takes the central window (what the evaluation wrappers do)."""
    random_crop_factor: Optional[float] = None
    resize_size: Tuple[int, int] = (180, 320)

    def window(self, h: int, w: int, rng: np.random.Generator, train: bool = True):
        if self.random_crop_factor is None:
            return 0, 0, h, w
        ch, cw = max(1, int(h * self.random_crop_factor)), max(1, int(w * self.random_crop_factor))
        if train:
            y0 = int(rng.integers(0, max(1, h - ch)))          # upper bound exclusive, like the reference's uniform draw
            x0 = int(rng.integers(0, max(1, w - cw)))
        else:
            y0, x0 = (h - ch) // 2, (w - cw) // 2
        return y0, x0, ch, cw

    def __call__(self, frames, rng: np.random.Generator, train: bool = True) -> torch.Tensor:
        frames = _decode_frames(frames)
        y0, x0, ch, cw = self.window(frames.shape[1], frames.shape[2], rng, train)
        x = frames[:, y0:y0 + ch, x0:x0 + cw].permute(0, 3, 1, 2).float() / 255.0
        x = F.interpolate(x, size=tuple(self.resize_size), mode="bilinear", align_corners=False)
        return x.permute(0, 2, 3, 1)


def rgb_to_hsv(x: torch.Tensor) -> torch.Tensor:
    """[..., 3] in [0, 1] → hue (turns, [0, 1)), saturation, value."""
    r, g, b = x.unbind(-1)
    v, _ = x.max(-1)
    c = v - x.min(-1)[0]
    safe = torch.where(c > 0, c, torch.ones_like(c))
    h = torch.where(v == r, ((g - b) / safe) % 6.0, torch.where(v == g, (b - r) / safe + 2.0, (r - g) / safe + 4.0))
    h = torch.where(c > 0, h / 6.0, torch.zeros_like(h))
    s = torch.where(v > 0, c / torch.where(v > 0, v, torch.ones_like(v)), torch.zeros_like(v))
    return torch.stack([h, s, v], -1)


def hsv_to_rgb(x: torch.Tensor) -> torch.Tensor:
    h, s, v = x.unbind(-1)
    k = (torch.tensor([5.0, 3.0, 1.0], device=x.device, dtype=x.dtype) + h.unsqueeze(-1) * 6.0) % 6.0
    return v.unsqueeze(-1) - (v * s).unsqueeze(-1) * torch.clamp(torch.minimum(k, 4.0 - k), 0.0, 1.0)


@dataclasses.dataclass(frozen=True)
class PhotometricDistortions:
    """Brightness, saturation, hue, contrast — in that order, each with ONE random level for the whole sequence
    (reference ``input_pipeline_rlds.py:390-452``; the ranges are its defaults).  Brightness adds a constant; saturation
    scales S of HSV (clipped to [0, 1]); hue rotates H by a fraction of a turn; contrast scales the distance to the
    per-frame, per-channel mean.  Like the TF ops the result is not clipped to [0, 1]."""
    brightness_max_delta: float = 0.1
    contrast_lower: float = 0.8
    contrast_upper: float = 1.2
    hue_max_delta: float = 0.03
    saturation_lower: float = 0.8
    saturation_upper: float = 1.2

    def sample(self, rng: np.random.Generator) -> Dict[str, float]:
        return {"brightness": float(rng.uniform(-self.brightness_max_delta, self.brightness_max_delta)),
                "saturation": float(rng.uniform(self.saturation_lower, self.saturation_upper)),
                "hue": float(rng.uniform(-self.hue_max_delta, self.hue_max_delta)),
                "contrast": float(rng.uniform(self.contrast_lower, self.contrast_upper))}

    def apply(self, frames: torch.Tensor, brightness=0.0, saturation=1.0, hue=0.0, contrast=1.0) -> torch.Tensor:
        x = frames
        if self.brightness_max_delta != 0.0:
            x = x + brightness
        change_s = self.saturation_lower != 1.0 or self.saturation_upper != 1.0
        change_h = self.hue_max_delta != 0.0
        if change_s or change_h:
            hsv = rgb_to_hsv(x)
            h, s, v = hsv.unbind(-1)
            if change_s:
                s = (s * saturation).clamp(0.0, 1.0)
            if change_h:
                h = (h + hue) % 1.0
            x = hsv_to_rgb(torch.stack([h, s, v], -1))
        if self.contrast_lower != 1.0 or self.contrast_upper != 1.0:
            mean = x.mean(dim=(-3, -2), keepdim=True)
            x = (x - mean) * contrast + mean
        return x

    def __call__(self, frames: torch.Tensor, rng: np.random.Generator) -> torch.Tensor:
        """frames float [t, h, w, 3] in [0, 1]."""
        return self.apply(frames, **self.sample(rng))


def random_resized_crop(frames: torch.Tensor, factor: Optional[float], out_hw, rng: np.random.Generator, train: bool) -> torch.Tensor:
    """Functional form of ``DecodeAndRandomResizedCrop``; values stay in the input's scale ([0, 255] for uint8 input)."""
    return DecodeAndRandomResizedCrop(factor, tuple(out_hw))(frames, rng, train) * 255.0


def photometric_distortions(frames: torch.Tensor, rng: np.random.Generator) -> torch.Tensor:
    """Functional form of ``PhotometricDistortions`` with the default ranges."""
    return PhotometricDistortions()(frames, rng)


class WindowDataset:
    """Iterable over (observation, action) windows of a list of episodes."""

    def __init__(self, episodes: List[List[Dict]], sequence_length: int, target_hw=(180, 320), random_crop_factor: Optional[float] = 0.95,
                 photometric: bool = True, train: bool = True, seed: int = 0, clip_tokenizer=None, clip_vocab_size: int = 49408):
        """``clip_tokenizer`` (a ``common.clip_tokenizer.ClipTokenizer``): episodes whose ``instruction`` is still the
        simulator's byte code (the "raw" tree of the converter) get ``instruction_tokenized_clip`` — [t, 77] CLIP BPE ids, what
        ``lava_lang_encoder="clip"`` feeds its text tower (reference ``input_pipeline_rlds.py:236-262``) — instead of
        ``clip_embedding``."""
        self.clip_tokenizer, self.clip_vocab_size, self._token_cache = clip_tokenizer, clip_vocab_size, {}
        self.windows = [w for ep in episodes for w in _windows(ep, sequence_length)]
        self.hw, self.factor, self.photometric, self.train = tuple(target_hw), random_crop_factor, photometric, train
        self.crop = DecodeAndRandomResizedCrop(random_crop_factor, self.hw)
        self.distort = PhotometricDistortions()
        self.rng = np.random.default_rng(seed)

    def __len__(self):
        return len(self.windows)

    def example(self, i: int) -> Dict:
        win = self.windows[i]
        first = win[0]["rgb"]
        encoded = isinstance(first, (bytes, bytearray, np.bytes_))
        rgb = self.crop([s["rgb"] for s in win] if encoded else np.stack([np.asarray(s["rgb"]) for s in win]), self.rng, self.train)
        if self.train and self.photometric:
            rgb = self.distort(rgb, self.rng)
        eff = torch.from_numpy(np.stack([np.asarray(s.get("effector_translation", np.zeros(2)), dtype=np.float32) for s in win]))
        obs = {"rgb": rgb, "effector_translation": eff}
        inst = np.asarray(win[0]["instruction"])
        if inst.dtype.kind in "iu" or inst.dtype.kind in "SU":       # byte-coded (or plain) text: tokens for the text tower
            if self.clip_tokenizer is None:
                raise ValueError("the episodes carry text instructions, not sentence embeddings: pass clip_tokenizer "
                                 "(config.clip_bpe_path) or convert them with an embedding (rlds_np_convert.py)")
            obs["instruction_tokenized_clip"] = torch.from_numpy(np.stack([self._tokens(s["instruction"]) for s in win]))
        else:
            obs["clip_embedding"] = torch.from_numpy(np.stack([np.asarray(s["instruction"], dtype=np.float32) for s in win]))
        return {"observation": obs, "action": torch.from_numpy(np.asarray(win[-1]["action"], dtype=np.float32))}

    def _tokens(self, instruction) -> np.ndarray:
        from language_table.common import clip_tokenizer as ct
        inst = np.asarray(instruction)
        text = str(inst) if inst.dtype.kind in "SU" else bytes(inst[inst != 0].astype(np.uint8).tolist()).decode("utf-8")
        if text not in self._token_cache:
            self._token_cache[text] = ct.tokenize_text(text, self.clip_tokenizer, vocab_size=self.clip_vocab_size)[0]
        return self._token_cache[text]

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
    tokenizer = None
    if getattr(config, "clip_bpe_path", None):                   # CLIP's merge table (not shipped; see common/clip_tokenizer.py)
        from language_table.common import clip_tokenizer as ct
        tokenizer = ct.ClipTokenizer(*ct.create_vocab(bpe_path=config.clip_bpe_path))
    ds = WindowDataset(eps, config.sequence_length, (config.data_target_height, config.data_target_width),
                       config.random_crop_factor if config.image_augment_crop else None, config.image_photometric_distortions,
                       train=True, seed=seed + process_index, clip_tokenizer=tokenizer)
    if normalization_path is not None and process_index != 0:
        return ds, _wait_for_norm_info(normalization_path)
    info = _load_norm_info(normalization_path) if normalization_path is not None else None
    if info is None:
        stat_batches = ((b["observation"], b["action"]) for b in ds.batches(min(8, len(ds)), shuffle=False, drop_last=False))
        obs_stats, act_stats, amin, amax = normalization.compute_dataset_statistics(
            (({k: v.numpy() for k, v in o.items()}, a.numpy()) for o, a in stat_batches), config.data_normalization_num_samples,
            skip_keys=tuple(config.data_skip_normalize_keys) + ("clip_embedding", "instruction_tokenized_clip"))
        info = {"observation_statistics": obs_stats, "action_statistics": act_stats, "action_min": amin, "action_max": amax}
        if normalization_path is not None:
            _save_norm_info(normalization_path, info)
    return ds, info
