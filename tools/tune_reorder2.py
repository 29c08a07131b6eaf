"""RT_TRACE_REORDER through the public call (keys + sort + indirect traversal, all timed)
against the plain call (development tool)."""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpu_raytracer_b200 as g  # noqa: E402
from gpu_raytracer_b200.configs import WORKLOADS  # noqa: E402
from tools.tune_trace import time_trace  # noqa: E402

for name in sys.argv[1:] or ["soup1m", "flake10m"]:
    wl = WORKLOADS[name]
    tris = g.gen_scene(*wl.scene)
    with g.Bvh.build(tris) as bvh:
        box = bvh.bounds
        prim = g.gen_primary_rays(wl.camera(), wl.width, wl.height)
        _, hits = time_trace(bvh, prim, 0, 1)
        radius = 0.05 * float(np.linalg.norm(box[3:] - box[:3]))
        sets = {"primary": (prim, 0),
                "ao_any": (g.gen_ao_rays(tris, prim, hits, 4, 2, radius), 1),
                "random4m": (g.gen_random_rays(box, 4, 0, 1 << 22), 0),
                "random32m": (g.gen_random_rays(box, 4, 0, 1 << 25), 0),
                "random100m": (g.gen_random_rays(box, 4, 0, 100_000_000), 0)}
        for kind, (rays, flags) in sets.items():
            ms0, ref = time_trace(bvh, rays, flags, 3)
            ms1, out = time_trace(bvh, rays, flags | g.RT_TRACE_REORDER, 3)
            same = bool(torch.equal(out.view(torch.int32), ref.view(torch.int32)))
            print(json.dumps(dict(scene=name, kind=kind, rays=rays.shape[0], plain_ms=round(ms0, 4),
                                  reorder_ms=round(ms1, 4), speedup=round(ms0 / ms1, 3), same=same)), flush=True)
            del ref, out
        del sets
    del tris
    torch.cuda.empty_cache()
