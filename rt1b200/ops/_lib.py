"""Loader for the native kernel library (``rt1b200/lib/librt1b200.so``).

The library is a plain C-ABI shared object built by ``nvcc`` for sm_100a from ``rt1b200/csrc/*.cu`` (see
``rt1b200/build.py``; declarations in ``include/rt1b200.h``).  It has no PyTorch or Python dependency: the
bindings below pass raw device pointers and the current CUDA stream through ``ctypes``.  Keeping the .so
in-tree (not in site-packages, not in a JIT cache) is deliberate: it travels with a repository snapshot.
"""
from __future__ import annotations

import ctypes
import os
from typing import Optional

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(_HERE), "lib", "librt1b200.so")

_lib: Optional[ctypes.CDLL] = None
_err: Optional[str] = None
_tried = False


def _try_load():
    global _lib, _err, _tried
    if _tried:
        return
    _tried = True
    if not os.path.exists(LIB_PATH):
        _err = f"{LIB_PATH} does not exist"
        return
    try:
        _lib = ctypes.CDLL(LIB_PATH, mode=ctypes.RTLD_GLOBAL)
        _lib.rt1_last_error.restype = ctypes.c_char_p
        _lib.rt1_abi_version.restype = ctypes.c_int
    except OSError as e:  # pragma: no cover - depends on the box
        _lib, _err = None, str(e)


def available() -> bool:
    _try_load()
    return _lib is not None


def load_error() -> Optional[str]:
    _try_load()
    return _err


def reload():
    """Forget a cached load attempt (used right after ``build()``)."""
    global _lib, _err, _tried
    _lib, _err, _tried = None, None, False
    _try_load()


def lib() -> ctypes.CDLL:
    _try_load()
    if _lib is None:
        raise RuntimeError(f"rt1b200 native library unavailable: {_err}")
    return _lib


_launches = 0


def launch_count() -> int:
    """Number of native kernel-launching calls made in this process: calls through :func:`call`, plus — for every replay
    of a captured step — the native calls that were recorded into its CUDA graph (:func:`add_launches`)."""
    return _launches


def add_launches(n: int) -> None:
    """Account for ``n`` native launches that a CUDA-graph replay just issued on the caller's behalf."""
    global _launches
    _launches += int(n)


def call(name: str, *args) -> None:
    """Invoke ``int name(...)`` from the library; raise with the library's message on a non-zero status."""
    global _launches
    fn = getattr(lib(), name)
    rc = fn(*args)
    _launches += 1
    if rc != 0:
        msg = lib().rt1_last_error()
        raise RuntimeError(f"{name} failed (status {rc}): {msg.decode() if msg else 'unknown'}")


# ctypes argument helpers -------------------------------------------------------------------------------
def ptr(t) -> ctypes.c_void_p:
    """Device/host pointer of a tensor (``None`` → NULL)."""
    return ctypes.c_void_p(0 if t is None else t.data_ptr())


def i32(v) -> ctypes.c_int:
    return ctypes.c_int(int(v))


def i64(v) -> ctypes.c_longlong:
    return ctypes.c_longlong(int(v))


def f32(v) -> ctypes.c_float:
    return ctypes.c_float(float(v))


def stream() -> ctypes.c_void_p:
    import torch
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
