"""The five workloads of BASELINE.json `configs` as concrete inputs
(SURVEY.md §8d): scene generator parameters, seeds, cameras, ray counts.

Host logic only; shared by bench.py, the tests and the tools so that every
measurement and every parity check talks about the same bytes.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional, Tuple

import numpy as np

from . import _lib as L
from . import camera


@dataclass(frozen=True)
class Workload:
    name: str
    scene: Tuple[int, int, int, int]        # kind, p0, p1, seed  (rt_gen_scene)
    n_tris: int
    key_bits: int
    width: int                               # primary-ray image; 0 = no primary rays
    height: int
    eye: Optional[Tuple[float, float, float]]
    look_at: Optional[Tuple[float, float, float]]
    vfov: float = 40.0
    ao_spp: int = 0                          # occlusion rays per pixel (any-hit)
    ao_radius_frac: float = 0.05             # of the scene diagonal (§8d config 3)
    ao_seed: int = 2
    random_rays: int = 0                     # config 5: incoherent rays, seed 4
    random_seed: int = 4
    note: str = ""

    def camera(self, width: Optional[int] = None, height: Optional[int] = None) -> np.ndarray:
        w, h = width or self.width, height or self.height
        return camera.pinhole(self.eye, self.look_at, (0.0, 1.0, 0.0), self.vfov, w, h)

    @property
    def primary_rays(self) -> int:
        return self.width * self.height


# sphereflake extent: root radius 1, children at 4/3 with radius 1/3, ... the
# whole flake fits in a ball of radius ~1.89 around the origin
_FLAKE_EYE = (1.7, 1.5, 4.2)

WORKLOADS = {
    # configs[0]: procedural Cornell box (~1K tris), 256x256 primary rays
    "cornell": Workload("cornell", (L.RT_SCENE_CORNELL, 0, 0, 0), 1024, 30, 256, 256,
                        (0.5, 0.5, -1.5), (0.5, 0.5, 0.5),
                        note="config 1: the CPU-runnable parity case"),
    # configs[1]: 1M-triangle random soup + 1920x1080 primary rays (the N=1 bench workload)
    "soup1m": Workload("soup1m", (L.RT_SCENE_SOUP, 1000000, 0, 1), 1000000, 30, 1920, 1080,
                       (0.5, 0.5, -0.9), (0.5, 0.5, 0.5),
                       note="config 2: the unit cube fills ~95 % of the 16:9 frame"),
    # configs[2]: 10M-triangle sphereflake, 4K primary + 4 AO rays per pixel
    "flake10m": Workload("flake10m", (L.RT_SCENE_SPHEREFLAKE, 4, 13, 2), 9979112, 30, 3840, 2160,
                         _FLAKE_EYE, (0.0, 0.0, 0.3), vfov=40.0, ao_spp=4,
                         note="config 3: depth 4 (7,381 spheres) x 8*13^2 triangles = 10M - 0.21 %"),
    # configs[3]: 100M-triangle terrain, 63-bit keys, build only
    "terrain100m": Workload("terrain100m", (L.RT_SCENE_TERRAIN, 7072, 0, 3), 100026368, 63, 0, 0,
                            None, None, note="config 4: build-scaling stress, 63-bit Morton"),
    # configs[4]: 1e9 incoherent random rays against the 10M scene, sharded over GPUs
    "random1b": Workload("random1b", (L.RT_SCENE_SPHEREFLAKE, 4, 13, 2), 9979112, 30, 0, 0,
                         None, None, random_rays=10**9,
                         note="config 5: rays generated on device per shard from (seed, index)"),
}
