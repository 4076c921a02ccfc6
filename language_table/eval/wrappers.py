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
