"""Profiler annotations.

The reference's RT-1 path has no tracing at all (its vendored JAX trainer uses ``jax.profiler.StepTraceAnnotation``,
``language_table/train/train.py:182``).  Here ``RT1B200_NVTX=1`` turns :func:`range_` into NVTX ranges (visible in Nsight
Systems / Compute timelines): training step → forward / backward / optimizer, every MBConv block of the image tokenizer, every
decoder layer.  Disabled (the default) it is a no-op context manager, so the hot path pays one attribute test per range —
and nothing at all inside a replayed CUDA graph.
"""
from __future__ import annotations

import contextlib
import os

ENABLED = os.environ.get("RT1B200_NVTX", "0") == "1"
_NULL = contextlib.nullcontext()


class _Range:
    __slots__ = ("name",)

    def __init__(self, name: str):
        self.name = name

    def __enter__(self):
        import torch
        torch.cuda.nvtx.range_push(self.name)

    def __exit__(self, *exc):
        import torch
        torch.cuda.nvtx.range_pop()


def range_(name: str):
    """``with trace.range_("block 3 fwd"): ...``"""
    return _Range(name) if ENABLED else _NULL


def enable(on: bool = True):
    global ENABLED
    ENABLED = bool(on)
