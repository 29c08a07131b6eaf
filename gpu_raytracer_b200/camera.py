"""Pinhole camera → the 12 floats rt_gen_primary_rays takes (A8).

Host logic only (float64 numpy, rounded once to float32), so the same camera
bytes go to the CUDA generator and to the oracle in the tests.
"""
from __future__ import annotations

import numpy as np


def pinhole(eye, look_at, up, vfov_deg: float, width: int, height: int) -> np.ndarray:
    """eye.xyz, corner.xyz, du.xyz, dv.xyz: pixel (x, y) looks through
    corner + (x + 0.5) du + (y + 0.5) dv; row 0 is the top of the image."""
    eye = np.asarray(eye, dtype=np.float64)
    fwd = np.asarray(look_at, dtype=np.float64) - eye
    fwd /= np.linalg.norm(fwd)
    right = np.cross(fwd, np.asarray(up, dtype=np.float64))
    right /= np.linalg.norm(right)
    upv = np.cross(right, fwd)
    half_h = np.tan(np.radians(vfov_deg) / 2.0)
    half_w = half_h * width / height
    corner = eye + fwd - right * half_w + upv * half_h
    du = right * (2.0 * half_w / width)
    dv = -upv * (2.0 * half_h / height)
    return np.concatenate([eye, corner, du, dv]).astype(np.float32)


def frame_box(box, width: int, height: int, vfov_deg: float = 40.0, direction=(0.3, 0.25, -1.0),
              fill: float = 1.0) -> np.ndarray:
    """A camera looking at the centre of an AABB from ``direction`` (pointing from
    the centre towards the eye), far enough that the box's bounding sphere fills
    ``fill`` of the vertical field of view."""
    box = np.asarray(box, dtype=np.float64)
    centre = 0.5 * (box[:3] + box[3:])
    radius = 0.5 * np.linalg.norm(box[3:] - box[:3])
    d = np.asarray(direction, dtype=np.float64)
    d /= np.linalg.norm(d)
    dist = radius / (np.sin(np.radians(vfov_deg) / 2.0) * fill)
    return pinhole(centre + d * dist, centre, (0.0, 1.0, 0.0), vfov_deg, width, height)


# the §8(d) camera of config 1: Cornell box seen from (0.5, 0.5, -1.5)
def cornell_camera(width: int = 256, height: int = 256) -> np.ndarray:
    return pinhole((0.5, 0.5, -1.5), (0.5, 0.5, 0.5), (0.0, 1.0, 0.0), 40.0, width, height)


def interlaced_field(cam, rank: int, world: int) -> np.ndarray:
    """Camera of ``rank``'s field of a frame with ``world`` times the rows of ``cam``'s.

    The tall frame (same frustum, ``height * world`` rows) is dealt to the ranks row by
    row: rank r renders rows r, r + world, r + 2 world, ...  Every field covers the whole
    frustum, so the ranks get statistically equal work, where contiguous bands give the
    ranks that see the middle of the scene several times the work of those that see its
    border ("contiguous ranges of the ray batch (or image tiles)").  The
    field is itself a ``width x height`` pinhole image: row y of the field looks through
    corner + (y * world + rank + 0.5) * dv / world, i.e. through the tall frame's row
    y * world + rank."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    c = np.asarray(cam, dtype=np.float64).reshape(12).copy()
    dv = c[9:12]
    c[3:6] = c[3:6] + dv * ((rank + 0.5) / world - 0.5)
    return c.astype(np.float32)
