"""Observation wrappers for closed-loop evaluation (reference ``language_table/eval/wrappers.py:26,64`` plus the
tf-agents ``HistoryWrapper`` it is combined with in ``main_rt1.py:130-142``), on plain ``reset/step`` environments.

  UseTokenWrapper          decode the byte instruction on reset, embed it once (512-d), attach it to every observation
  CentralCropImageWrapper  float convert, 95 % centre crop, bilinear resize to (height, width) → ``rgb_sequence``
  HistoryWrapper           keep the last ``history_length`` observations, tiling the first one on reset
"""
import collections

import numpy as np
import torch

from language_table.common import rt1_tokenizer
from language_table.environments.language_table import LanguageTable
from rt1b200.envs import raster


class _Wrapper:
    def __init__(self, env):
        self._env = env

    def __getattr__(self, name):
        return getattr(self._env, name)

    def reset(self):
        return self._process(self._env.reset(), reset=True)

    def step(self, action):
        obs, reward, done, info = self._env.step(action)
        return self._process(obs, reset=False), reward, done, info


class UseTokenWrapper(_Wrapper):
    def __init__(self, env):
        super().__init__(env)
        self._current_tokens = None

    def _process(self, obs, reset):
        if reset or self._current_tokens is None:
            text = LanguageTable.decode_instruction(obs["instruction"])
            self._current_tokens = rt1_tokenizer.tokenize_text(text)
        obs = collections.OrderedDict(obs)
        obs["instruction_tokenized_use"] = self._current_tokens
        return obs


class ClipTokenWrapper(_Wrapper):
    """Adds ``instruction_tokenized_clip`` — int64 [77] CLIP BPE ids of the decoded instruction, framed by the start / end tokens —
    what a LAVA model with ``lava_lang_encoder="clip"`` feeds its text tower.  Upstream Language-Table's evaluation script asks
    for this wrapper (reference ``eval/main.py:85``); the fork's ``wrappers.py`` only carries ``UseTokenWrapper``.  ``tokenizer``: a
    ``common.clip_tokenizer.ClipTokenizer``, or ``bpe_path`` to CLIP's merge file (not shipped)."""

    def __init__(self, env, tokenizer=None, bpe_path=None, vocab_size: int = 49408, context_length: int = 77):
        super().__init__(env)
        from language_table.common import clip_tokenizer
        if tokenizer is None:
            if not bpe_path:
                raise ValueError("ClipTokenWrapper needs a tokenizer or the path of CLIP's BPE merge file (bpe_simple_vocab_16e6.txt.gz)")
            tokenizer = clip_tokenizer.ClipTokenizer(*clip_tokenizer.create_vocab(bpe_path=bpe_path))
        self._tokenize = lambda text: clip_tokenizer.tokenize_text(text, tokenizer, vocab_size=vocab_size, context_length=context_length)[0]
        self._current_tokens = None

    def _process(self, obs, reset):
        if reset or self._current_tokens is None:
            self._current_tokens = self._tokenize(LanguageTable.decode_instruction(obs["instruction"]))
        obs = collections.OrderedDict(obs)
        obs["instruction_tokenized_clip"] = self._current_tokens
        return obs


def crop_test_image(images, random_crop_factor):
    """The centre crop that is the mean of the training-time random crop (reference ``wrappers.py:99-123``): a window of
    ``int(H * f) x int(W * f)`` at offset ``((H - H f) // 2, (W - W f) // 2)``.  ``images``: ``[H, W, C]`` or ``[N, H, W, C]``
    (numpy or torch)."""
    h, w = images.shape[-3], images.shape[-2]
    sh, sw = h * random_crop_factor, w * random_crop_factor
    oy, ox = int((h - sh) // 2), int((w - sw) // 2)
    return images[..., oy:oy + int(sh), ox:ox + int(sw), :]


def resize_images(images, target_height=None, target_width=None):
    """Bilinear resize (half-pixel centres, no antialiasing — ``tf.image.resize``'s default, reference ``:126-137``) of
    ``[H, W, C]`` or ``[N, H, W, C]`` images to float32 ``[..., target_height, target_width, C]``."""
    assert target_height and target_width
    x = torch.as_tensor(np.asarray(images)).float()
    single = x.dim() == 3
    x = (x[None] if single else x).permute(0, 3, 1, 2)
    y = torch.nn.functional.interpolate(x, size=(target_height, target_width), mode="bilinear", align_corners=False, antialias=False)
    y = y.permute(0, 2, 3, 1)
    return (y[0] if single else y).numpy()


class CentralCropImageWrapper(_Wrapper):
    def __init__(self, env, target_height=256, target_width=456, random_crop_factor=0.95):
        super().__init__(env)
        self._h, self._w, self._crop = target_height, target_width, random_crop_factor

    def _process(self, obs, reset):
        obs = collections.OrderedDict(obs)
        frame = torch.from_numpy(np.ascontiguousarray(obs["rgb"]))[None]
        img = raster.crop_resize(frame, self._h, self._w, self._crop)[0]          # [3,H,W] float in [0,1]
        obs["rgb_sequence"] = img.permute(1, 2, 0).numpy()
        return obs


class HistoryWrapper(_Wrapper):
    def __init__(self, env, history_length=6, tile_first_step_obs=True):
        super().__init__(env)
        self._n, self._tile = history_length, tile_first_step_obs
        self._buf = collections.deque(maxlen=history_length)

    def _process(self, obs, reset):
        if reset:
            self._buf.clear()
            for _ in range(self._n if self._tile else 1):
                self._buf.append(obs)
        else:
            self._buf.append(obs)
        frames = list(self._buf)
        while len(frames) < self._n:
            frames.insert(0, {k: np.zeros_like(v) for k, v in frames[0].items()})
        return collections.OrderedDict((k, np.stack([np.asarray(f[k]) for f in frames])) for k in frames[0])
