"""Dataset statistics and (de)normalisers for the LAVA baseline (reference ``language_table/train/normalization.py``).

``ChanRunningStatistics`` merges batches of samples into a running mean / variance with Chan et al.'s parallel update;
``compute_dataset_statistics`` makes one pass over ``num_samples`` batches and returns per-key observation statistics,
action statistics and the per-dimension action range; the callable normalisers mirror the reference's small classes.
"""
from typing import Dict, Iterable, Optional, Sequence, Tuple

import numpy as np

EPS = np.finfo(np.float32).eps


class ChanRunningStatistics:
    def __init__(self, initial_mean: Optional[np.ndarray] = None):
        self._n = 0
        self._mean = None if initial_mean is None else np.zeros_like(np.asarray(initial_mean, dtype=np.float64))
        self._m2 = None if initial_mean is None else np.zeros_like(self._mean)

    def update_running_statistics(self, sample: np.ndarray):
        """``sample``: one vector [d] or a batch [n, d]."""
        s = np.asarray(sample, dtype=np.float64)
        if s.ndim == 1:
            s = s[None]
        nb, mb = s.shape[0], s.mean(axis=0)
        m2b = ((s - mb) ** 2).sum(axis=0)
        if self._n == 0:
            self._n, self._mean, self._m2 = nb, mb, m2b
            return
        n = self._n + nb
        delta = mb - self._mean
        self._mean = self._mean + delta * nb / n
        self._m2 = self._m2 + m2b + delta ** 2 * self._n * nb / n
        self._n = n

    @property
    def mean(self):
        return self._mean

    @property
    def variance(self):
        return self._m2 / max(self._n, 1)

    @property
    def std(self):
        return np.sqrt(self.variance)

    @property
    def n(self):
        return self._n


def compute_dataset_statistics(batches: Iterable[Tuple[Dict[str, np.ndarray], np.ndarray]], num_samples: int,
                               skip_keys: Sequence[str] = ("rgb", "instruction")):
    """``batches`` yields ``(observation dict, action [b, ..., A])``.  Returns ``(obs_stats, act_stats, min_action,
    max_action)`` with ``*_stats = {"mean": ..., "std": ...}`` (per key for observations)."""
    obs_stats: Dict[str, ChanRunningStatistics] = {}
    act = ChanRunningStatistics()
    amin = amax = None
    for i, (obs, action) in enumerate(batches):
        if i >= num_samples:
            break
        a = np.asarray(action, dtype=np.float64).reshape(-1, np.asarray(action).shape[-1])
        act.update_running_statistics(a)
        amin = a.min(axis=0) if amin is None else np.minimum(amin, a.min(axis=0))
        amax = a.max(axis=0) if amax is None else np.maximum(amax, a.max(axis=0))
        for k, v in obs.items():
            if k in skip_keys:
                continue
            v = np.asarray(v, dtype=np.float64)
            obs_stats.setdefault(k, ChanRunningStatistics()).update_running_statistics(v.reshape(-1, v.shape[-1]))
    pack = lambda s: {"mean": s.mean.astype(np.float32), "std": np.maximum(s.std, EPS).astype(np.float32)}    # noqa: E731
    return ({k: pack(s) for k, s in obs_stats.items()}, pack(act), amin.astype(np.float32), amax.astype(np.float32))


class IdentityNormalizer:
    def __init__(self, cast_dtype=np.float32):
        self.cast_dtype = cast_dtype

    def __call__(self, x, **kwargs):
        return np.asarray(x).astype(self.cast_dtype)


class StdNormalizer:
    def __init__(self, mean, std):
        self.mean, self.std = np.asarray(mean, np.float32), np.maximum(np.asarray(std, np.float32), EPS)

    def __call__(self, vector, **kwargs):
        return (vector - self.mean) / self.std


class StdDenormalizer(StdNormalizer):
    def __call__(self, vector, mean_offset=True, **kwargs):
        out = vector * self.std
        return out + self.mean if mean_offset else out


class MinMaxNormalizerBase:
    """Centre and half-width of ``[vmin, vmax]``; the half-width is floored at ``EPS`` so a constant feature does not
    divide by zero (reference ``normalization.py:306-314``)."""

    def __init__(self, vmin, vmax):
        self.vmin, self.vmax = np.asarray(vmin, np.float32), np.asarray(vmax, np.float32)
        self.mean_range = (self.vmin + self.vmax) / 2.0
        self.half_range = np.maximum(0.5 * (self.vmax - self.vmin), EPS)


class MinMaxNormalizer(MinMaxNormalizerBase):
    """Maps [vmin, vmax] to [-1, 1]."""

    def __call__(self, vector, **kwargs):
        return (vector - self.mean_range) / self.half_range


class MinMaxDenormalizer(MinMaxNormalizerBase):
    """Maps [-1, 1] back to [vmin, vmax]."""

    def __call__(self, vector, **kwargs):
        return vector * self.half_range + self.mean_range


def construct_normalizers(obs_stats: Dict, act_stats: Dict, skip_keys: Sequence[str] = ("rgb", "instruction")):
    """Per-key observation normalisers, the action normaliser and the action denormaliser."""
    obs_norm = {k: StdNormalizer(v["mean"], v["std"]) for k, v in obs_stats.items() if k not in skip_keys}
    return obs_norm, StdNormalizer(act_stats["mean"], act_stats["std"]), StdDenormalizer(act_stats["mean"], act_stats["std"])
