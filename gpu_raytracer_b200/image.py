"""Image output (SURVEY.md §8f.1 "PPM/PNG writer"): float RGB → binary PPM or PNG; numpy and the
standard library only."""
from __future__ import annotations

import struct
import zlib

import numpy as np

# 8-bit code of a linear value c in [0, 1] with display gamma 2.2: round(c^(1/2.2) * 255).
# Written as a search in 255 thresholds, t_v = ((v - 0.5) / 255)^2.2, so that the code of a
# float32 input does not depend on the platform's powf: code = number of thresholds <= c.
_THRESHOLDS = ((np.arange(1, 256, dtype=np.float64) - 0.5) / 255.0) ** 2.2


def quantize(rgb) -> np.ndarray:
    """uint8 codes of linear float values (any shape); NaN and negatives give 0."""
    a = np.nan_to_num(np.asarray(rgb, dtype=np.float32), nan=0.0).astype(np.float64)
    return np.searchsorted(_THRESHOLDS, np.clip(a, 0.0, 1.0), side="right").astype(np.uint8)


def write_ppm(path: str, rgb, width: int, height: int) -> None:
    """rgb: [height*width, 3] floats in [0, 1] (numpy or torch, any device)."""
    if hasattr(rgb, "detach"):
        rgb = rgb.detach().cpu().numpy()
    a = quantize(np.asarray(rgb, dtype=np.float32).reshape(height, width, 3))
    with open(path, "wb") as f:
        f.write(f"P6\n{width} {height}\n255\n".encode())
        f.write(a.tobytes())


def write_png(path: str, rgb, width: int, height: int, level: int = 6) -> None:
    """The same pixels as write_ppm in an 8-bit RGB PNG (filter type 0 rows, one IDAT chunk)."""
    if hasattr(rgb, "detach"):
        rgb = rgb.detach().cpu().numpy()
    a = quantize(np.asarray(rgb, dtype=np.float32).reshape(height, width, 3))
    rows = np.concatenate([np.zeros((height, 1), dtype=np.uint8), a.reshape(height, width * 3)], axis=1)

    def chunk(tag: bytes, data: bytes) -> bytes:
        return struct.pack(">I", len(data)) + tag + data + struct.pack(">I", zlib.crc32(tag + data) & 0xFFFFFFFF)

    with open(path, "wb") as f:
        f.write(b"\x89PNG\r\n\x1a\n")
        f.write(chunk(b"IHDR", struct.pack(">IIBBBBB", width, height, 8, 2, 0, 0, 0)))
        f.write(chunk(b"IDAT", zlib.compress(rows.tobytes(), level)))
        f.write(chunk(b"IEND", b""))
