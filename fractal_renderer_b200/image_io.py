"""Lossless image output for render artefacts.  The reference writes AVIF through ravif/rav1e
(src/lib.rs:323-345), which is unavailable here (needs cargo + NASM) and lossy; PPM and PNG are written instead, with
nothing but the standard library, so that frames can be diffed bit for bit."""
from __future__ import annotations

import struct
import zlib

import numpy as np


def _as_rgb(image: np.ndarray, width: int = 0, height: int = 0) -> np.ndarray:
    """Accepts get_image's (H*W, 3) Vec<RGB> image, (H, W, 3) or (H, W, 4) arrays."""
    a = np.asarray(image, dtype=np.uint8)
    if a.ndim == 2 and a.shape[1] == 3:
        if width * height != a.shape[0]:
            raise ValueError("width*height must equal the number of pixels")
        a = a.reshape(height, width, 3)
    if a.ndim != 3 or a.shape[2] not in (3, 4):
        raise ValueError("expected (H*W,3), (H,W,3) or (H,W,4) uint8")
    return a


def write_ppm(path: str, image: np.ndarray, width: int = 0, height: int = 0) -> None:
    a = _as_rgb(image, width, height)[..., :3]
    with open(path, "wb") as f:
        f.write(f"P6\n{a.shape[1]} {a.shape[0]}\n255\n".encode())
        f.write(np.ascontiguousarray(a).tobytes())


def read_ppm(path: str) -> np.ndarray:
    data = open(path, "rb").read()
    parts = data.split(b"\n", 3)
    if parts[0] != b"P6" or parts[2] != b"255":
        raise ValueError("not a binary 8-bit PPM")
    w, h = (int(v) for v in parts[1].split())
    return np.frombuffer(parts[3], np.uint8, w * h * 3).reshape(h, w, 3).copy()


def _chunk(tag: bytes, payload: bytes) -> bytes:
    return struct.pack(">I", len(payload)) + tag + payload + struct.pack(">I", zlib.crc32(tag + payload) & 0xFFFFFFFF)


def write_png(path: str, image: np.ndarray, width: int = 0, height: int = 0, level: int = 6) -> None:
    a = np.ascontiguousarray(_as_rgb(image, width, height))
    h, w, c = a.shape
    raw = np.empty((h, 1 + w * c), np.uint8)
    raw[:, 0] = 0  # filter type None
    raw[:, 1:] = a.reshape(h, w * c)
    ihdr = struct.pack(">IIBBBBB", w, h, 8, 2 if c == 3 else 6, 0, 0, 0)
    with open(path, "wb") as f:
        f.write(b"\x89PNG\r\n\x1a\n" + _chunk(b"IHDR", ihdr) + _chunk(b"IDAT", zlib.compress(raw.tobytes(), level)) + _chunk(b"IEND", b""))
