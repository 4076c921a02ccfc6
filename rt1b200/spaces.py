"""Minimal, dependency-free tensor-space descriptions.

The reference describes observations, actions and the rolling network state with
``gym.spaces`` objects (``distribute_train.py:28-40``,
``pytorch_robotics_transformer/transformer_network.py:105-123``).  ``gym`` is not a
dependency of this framework: the four space kinds RT-1 needs are re-stated here
with the same attribute names (``shape``, ``low``, ``high``, ``n``, ``nvec``,
``sample()``, ``keys()``/``__getitem__`` for ``Dict``) so user code that only reads
those attributes keeps working.  ``Dict`` sorts plain-``dict`` input by key and
keeps ``OrderedDict`` order, which is what gym 0.23 did and what the action
tokenizer's ordering relies on (``action_tokenizer.py:82``).
"""
from __future__ import annotations

from collections import OrderedDict
from typing import Iterable, Mapping, Optional, Sequence, Union

import numpy as np

__all__ = ["Space", "Box", "Discrete", "MultiDiscrete", "Dict"]


class Space:
    shape: tuple = ()
    dtype = None

    def __init__(self, seed: Optional[int] = None):
        self._rng = np.random.default_rng(seed)

    def seed(self, seed: Optional[int] = None):
        self._rng = np.random.default_rng(seed)
        return [seed]

    def sample(self):
        raise NotImplementedError

    def contains(self, x) -> bool:
        raise NotImplementedError

    def __contains__(self, x) -> bool:
        return self.contains(x)


class Box(Space):
    """Axis-aligned box; ``low``/``high`` are broadcast to ``shape``."""

    def __init__(self, low, high, shape: Optional[Sequence[int]] = None,
                 dtype=np.float32, seed: Optional[int] = None):
        super().__init__(seed)
        self.dtype = np.dtype(dtype)
        if shape is None:
            shape = np.broadcast(np.asarray(low), np.asarray(high)).shape
        self.shape = tuple(int(s) for s in shape)
        self.low = np.broadcast_to(np.asarray(low, dtype=np.float64), self.shape).astype(self.dtype)
        self.high = np.broadcast_to(np.asarray(high, dtype=np.float64), self.shape).astype(self.dtype)

    def sample(self):
        lo = np.where(np.isfinite(self.low), self.low, 0.0).astype(np.float64)
        hi = np.where(np.isfinite(self.high), self.high, 0.0).astype(np.float64)
        bounded = np.isfinite(self.low) & np.isfinite(self.high)
        out = np.where(bounded, self._rng.uniform(lo, np.maximum(hi, lo)),
                       self._rng.standard_normal(self.shape))
        return out.astype(self.dtype)

    def contains(self, x) -> bool:
        x = np.asarray(x)
        return x.shape == self.shape and bool(np.all(x >= self.low) and np.all(x <= self.high))

    def __repr__(self):
        return f"Box({self.low.min()}, {self.high.max()}, {self.shape}, {self.dtype})"


class Discrete(Space):
    def __init__(self, n: int, seed: Optional[int] = None):
        super().__init__(seed)
        self.n = int(n)
        self.shape = ()
        self.dtype = np.dtype(np.int64)

    def sample(self):
        return int(self._rng.integers(self.n))

    def contains(self, x) -> bool:
        return 0 <= int(x) < self.n

    def __repr__(self):
        return f"Discrete({self.n})"


class MultiDiscrete(Space):
    def __init__(self, nvec, dtype=np.int64, seed: Optional[int] = None):
        super().__init__(seed)
        self.nvec = np.asarray(nvec, dtype=dtype)
        self.shape = self.nvec.shape
        self.dtype = np.dtype(dtype)

    def sample(self):
        return (self._rng.random(self.nvec.shape) * self.nvec).astype(self.dtype)

    def contains(self, x) -> bool:
        x = np.asarray(x)
        return x.shape == self.shape and bool(np.all(x >= 0) and np.all(x < self.nvec))

    def __repr__(self):
        return f"MultiDiscrete({self.nvec.tolist()})"


class Dict(Space):
    def __init__(self, spaces: Union[Mapping[str, Space], Iterable, None] = None,
                 seed: Optional[int] = None, **kw):
        super().__init__(seed)
        if spaces is None:
            spaces = kw
        if isinstance(spaces, OrderedDict):
            items = list(spaces.items())
        elif isinstance(spaces, Mapping):
            items = sorted(spaces.items(), key=lambda kv: kv[0])
        else:
            items = list(spaces)
        self.spaces = OrderedDict(items)

    def keys(self):
        return self.spaces.keys()

    def values(self):
        return self.spaces.values()

    def items(self):
        return self.spaces.items()

    def __getitem__(self, k):
        return self.spaces[k]

    def __iter__(self):
        return iter(self.spaces)

    def __len__(self):
        return len(self.spaces)

    def sample(self):
        return OrderedDict((k, s.sample()) for k, s in self.spaces.items())

    def contains(self, x) -> bool:
        return all(k in x and s.contains(x[k]) for k, s in self.spaces.items())

    def __repr__(self):
        return "Dict(" + ", ".join(f"{k}: {v}" for k, v in self.spaces.items()) + ")"
