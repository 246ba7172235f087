"""Minimal image writers for headless output (no window / swapchain on this path)."""
from __future__ import annotations

import struct
import zlib

import numpy as np


def tonemap(rgb: np.ndarray) -> np.ndarray:
    """Simple Reinhard + gamma for quick previews; the reference's display transform (Tony McMapface LUT) is prt_tone_map / prosper_b200.tonemap."""
    c = np.nan_to_num(np.maximum(rgb, 0.0))
    return np.clip(c / (1.0 + c), 0, 1) ** (1 / 2.2)


def write_png(path: str, rgb8: np.ndarray) -> None:
    h, w, _ = rgb8.shape
    raw = b"".join(b"\x00" + rgb8[y].tobytes() for y in range(h))

    def chunk(tag, data):
        c = struct.pack(">I", len(data)) + tag + data
        return c + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)

    with open(path, "wb") as f:
        f.write(b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 2, 0, 0, 0))
                + chunk(b"IDAT", zlib.compress(raw, 6)) + chunk(b"IEND", b""))


def save_preview(path: str, rgba: np.ndarray) -> None:
    write_png(path, (tonemap(rgba[..., :3]) * 255 + 0.5).astype(np.uint8))
