"""Instruction → 512-d sentence embedding (reference ``language_table/common/rt1_tokenizer.py``: Universal Sentence
Encoder through TF-Hub, re-loaded on every call).  USE cannot be downloaded here, so the default encoder is a
deterministic stand-in: a hashed bag of word n-grams projected with a fixed random matrix and L2-normalised — stable
across processes, similar instructions get similar vectors.  Plug a real encoder in with :func:`set_encoder`."""
import hashlib
from typing import Callable, Optional

import numpy as np

_ENCODER: Optional[Callable[[str], np.ndarray]] = None
_DIM = 512


def set_encoder(fn: Optional[Callable[[str], np.ndarray]]):
    global _ENCODER
    _ENCODER = fn


def _hash_embed(text: str) -> np.ndarray:
    words = text.lower().split()
    grams = words + [" ".join(words[i:i + 2]) for i in range(len(words) - 1)]
    vec = np.zeros(_DIM, np.float64)
    for g in grams:
        h = hashlib.sha256(g.encode("utf-8")).digest()
        rng = np.random.RandomState(int.from_bytes(h[:4], "little"))
        vec += rng.standard_normal(_DIM)
    n = np.linalg.norm(vec)
    return (vec / n if n > 0 else vec).astype(np.float32)


def tokenize_text(text: str) -> np.ndarray:
    """512-d float32 embedding of an instruction."""
    return np.asarray((_ENCODER or _hash_embed)(text), np.float32).reshape(_DIM)
