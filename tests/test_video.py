"""Codec-free episode videos (reference language_table/eval/main_rt1.py:196-199 writes mp4 through imageio)."""
import struct
import zlib

import numpy as np

from rt1b200.utils import video


def _frames(t=5, h=18, w=32):
    rng = np.random.RandomState(0)
    base = rng.randint(0, 256, (h, w, 3)).astype(np.uint8)
    return np.stack([np.roll(base, i, axis=1) for i in range(t)])


def test_apng_round_trip(tmp_path):
    fr = _frames()
    path = video.write_apng(str(tmp_path / "ep.png"), fr, fps=10)
    data = open(path, "rb").read()
    assert data[:8] == b"\x89PNG\r\n\x1a\n"
    pos, chunks = 8, []
    while pos < len(data):
        n, tag = struct.unpack(">I4s", data[pos:pos + 8])
        body = data[pos + 8:pos + 8 + n]
        assert struct.unpack(">I", data[pos + 8 + n:pos + 12 + n])[0] == zlib.crc32(tag + body) & 0xFFFFFFFF
        chunks.append((tag, body))
        pos += 12 + n
    tags = [t for t, _ in chunks]
    assert tags[0] == b"IHDR" and tags[-1] == b"IEND" and tags.count(b"fcTL") == 5 and tags.count(b"fdAT") == 4
    assert struct.unpack(">II", dict(chunks)[b"acTL"]) == (5, 0)
    # decode every frame (filter 0 rows) and compare bit for bit
    h, w = fr.shape[1:3]
    got = []
    for tag, body in chunks:
        if tag in (b"IDAT", b"fdAT"):
            raw = zlib.decompress(body if tag == b"IDAT" else body[4:])
            rows = np.frombuffer(raw, np.uint8).reshape(h, 1 + 3 * w)
            assert (rows[:, 0] == 0).all()
            got.append(rows[:, 1:].reshape(h, w, 3))
    assert np.array_equal(np.stack(got), fr)
    # a still image viewer (PIL) reads the first frame
    from PIL import Image
    assert np.array_equal(np.asarray(Image.open(path).convert("RGB")), fr[0])


def test_gif_decodes_with_pil(tmp_path):
    from PIL import Image
    fr = _frames(t=4)
    path = video.write_gif(str(tmp_path / "ep.gif"), fr, fps=10)
    im = Image.open(path)
    assert im.n_frames == 4 and im.size == (32, 18)
    for i in range(4):
        im.seek(i)
        got = np.asarray(im.convert("RGB")).astype(int)
        assert np.abs(got - fr[i].astype(int)).max() <= 26          # 6x7x6 colour cube: at most half a palette step
    assert video.write_video(str(tmp_path / "x.mp4"), fr).endswith("x.png")
