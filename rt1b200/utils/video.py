"""Episode videos without a codec dependency.

The reference's evaluator writes every episode as an mp4 through imageio/ffmpeg (``language_table/eval/main_rt1.py:196-199``);
neither exists offline.  Two self-contained writers take their place — both play in any browser:

* :func:`write_apng` — animated PNG, lossless 24-bit colour (``zlib`` from the standard library does the compression);
* :func:`write_gif`  — animated GIF, a fixed 6x7x6 colour cube with LZW coding written here (for viewers without APNG).

``frames`` is ``uint8 [T, H, W, 3]`` (numpy array or anything convertible).
"""
from __future__ import annotations

import struct
import zlib
from typing import Sequence

import numpy as np


def _frames(frames) -> np.ndarray:
    arr = np.asarray(frames)
    if arr.dtype != np.uint8:
        arr = np.clip(arr, 0, 255).astype(np.uint8)
    if arr.ndim != 4 or arr.shape[-1] != 3 or arr.shape[0] == 0:
        raise ValueError(f"frames must be uint8 [T, H, W, 3] with T >= 1, got {arr.shape}")
    return np.ascontiguousarray(arr)


# ------------------------------------------------------------------------------------------------------ APNG
def _chunk(tag: bytes, data: bytes) -> bytes:
    return struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)


def write_apng(path: str, frames, fps: float = 10.0, loop: int = 0, level: int = 6) -> str:
    """Animated PNG: one IDAT frame followed by ``fdAT`` frames, every frame the full image (dispose none, blend source)."""
    arr = _frames(frames)
    t, h, w, _ = arr.shape
    num, den = 100, max(1, int(round(100 * fps)))               # frame delay = num / den seconds
    out = [b"\x89PNG\r\n\x1a\n", _chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 2, 0, 0, 0)),
           _chunk(b"acTL", struct.pack(">II", t, loop))]
    seq = 0
    for i in range(t):
        out.append(_chunk(b"fcTL", struct.pack(">IIIIIHHBB", seq, w, h, 0, 0, num, den, 0, 0)))
        seq += 1
        raw = np.concatenate([np.zeros((h, 1), np.uint8), arr[i].reshape(h, w * 3)], 1).tobytes()     # filter type 0 per row
        comp = zlib.compress(raw, level)
        if i == 0:
            out.append(_chunk(b"IDAT", comp))
        else:
            out.append(_chunk(b"fdAT", struct.pack(">I", seq) + comp))
            seq += 1
    out.append(_chunk(b"IEND", b""))
    with open(path, "wb") as f:
        f.write(b"".join(out))
    return path


# ------------------------------------------------------------------------------------------------------- GIF
_LEVELS = (6, 7, 6)             # red, green, blue steps of the fixed palette (252 colours)


def _palette() -> bytes:
    r, g, b = (np.round(np.linspace(0, 255, n)).astype(np.uint8) for n in _LEVELS)
    pal = np.stack(np.meshgrid(r, g, b, indexing="ij"), -1).reshape(-1, 3)
    pal = np.concatenate([pal, np.zeros((256 - len(pal), 3), np.uint8)], 0)
    return pal.tobytes()


def _quantise(frame: np.ndarray) -> np.ndarray:
    idx = [np.round(frame[..., c].astype(np.float32) * ((_LEVELS[c] - 1) / 255.0)).astype(np.int32) for c in range(3)]
    return ((idx[0] * _LEVELS[1] + idx[1]) * _LEVELS[2] + idx[2]).astype(np.uint8)


def _lzw(indices: Sequence[int], min_code_size: int = 8) -> bytes:
    """GIF-flavoured LZW: variable-width codes (LSB first), clear code at the start and whenever the table fills."""
    clear, eoi = 1 << min_code_size, (1 << min_code_size) + 1
    table = {bytes([i]): i for i in range(clear)}
    nxt, width = eoi + 1, min_code_size + 1
    acc, nbits, out = 0, 0, bytearray()

    def emit(code: int):
        nonlocal acc, nbits
        acc |= code << nbits
        nbits += width
        while nbits >= 8:
            out.append(acc & 0xFF)
            acc >>= 8
            nbits -= 8

    emit(clear)
    cur = b""
    for v in indices:
        ext = cur + bytes([v])
        if ext in table:
            cur = ext
            continue
        emit(table[cur])
        if nxt < 4096:
            table[ext] = nxt
            nxt += 1
            if nxt > (1 << width) and width < 12:
                width += 1
        else:                                                   # table full: start over
            emit(clear)
            table = {bytes([i]): i for i in range(clear)}
            nxt, width = eoi + 1, min_code_size + 1
        cur = bytes([v])
    if cur:
        emit(table[cur])
    emit(eoi)
    if nbits:
        out.append(acc & 0xFF)
    return bytes(out)


def write_gif(path: str, frames, fps: float = 10.0, loop: int = 0) -> str:
    arr = _frames(frames)
    t, h, w, _ = arr.shape
    delay = max(1, int(round(100.0 / fps)))                     # hundredths of a second
    out = [b"GIF89a", struct.pack("<HHBBB", w, h, 0xF7, 0, 0), _palette(),
           b"\x21\xFF\x0BNETSCAPE2.0\x03\x01" + struct.pack("<H", loop) + b"\x00"]
    for i in range(t):
        out.append(b"\x21\xF9\x04\x04" + struct.pack("<H", delay) + b"\x00\x00")
        out.append(b"\x2C" + struct.pack("<HHHHB", 0, 0, w, h, 0))
        data = _lzw(_quantise(arr[i]).reshape(-1).tolist())
        out.append(b"\x08")
        for k in range(0, len(data), 255):
            blk = data[k:k + 255]
            out.append(bytes([len(blk)]) + blk)
        out.append(b"\x00")
    out.append(b"\x3B")
    with open(path, "wb") as f:
        f.write(b"".join(out))
    return path


def write_video(path: str, frames, fps: float = 10.0) -> str:
    """Pick the writer from the extension (``.gif`` → GIF, anything else → APNG with a ``.png`` suffix)."""
    if path.lower().endswith(".gif"):
        return write_gif(path, frames, fps)
    if not path.lower().endswith((".png", ".apng")):
        path = path.rsplit(".", 1)[0] + ".png" if "." in path.rsplit("/", 1)[-1] else path + ".png"
    return write_apng(path, frames, fps)
