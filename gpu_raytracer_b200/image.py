"""Image output: float RGB → binary PPM, no dependencies."""
from __future__ import annotations

import numpy as np


def write_ppm(path: str, rgb, width: int, height: int) -> None:
    """rgb: [height*width, 3] floats in [0, 1] (numpy or torch, any device)."""
    if hasattr(rgb, "detach"):
        rgb = rgb.detach().cpu().numpy()
    a = np.clip(np.asarray(rgb, dtype=np.float32).reshape(height, width, 3), 0.0, 1.0)
    a = (a ** (1.0 / 2.2) * 255.0 + 0.5).astype(np.uint8)
    with open(path, "wb") as f:
        f.write(f"P6\n{width} {height}\n255\n".encode())
        f.write(a.tobytes())
