"""Upper bound of ray reordering (development tool): rays sorted by a coherence key with
torch, the permuted copy traced with the normal kernel; only the trace is timed."""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpu_raytracer_b200 as g  # noqa: E402
from gpu_raytracer_b200.configs import WORKLOADS  # noqa: E402
from tools.tune_trace import time_trace  # noqa: E402


def spread3(x):  # 10 bits -> every third bit
    x = x.long()
    out = torch.zeros_like(x)
    for b in range(10):
        out |= ((x >> b) & 1) << (3 * b)
    return out


def keys(rays, box, mode):
    lo = torch.tensor(box[:3], device="cuda"); ext = torch.tensor(box[3:] - box[:3], device="cuda")
    o = ((rays[:, 0:3] - lo) / ext).clamp(0, 0.999999)
    q = (o * 1024).long()
    mort = (spread3(q[:, 0]) << 2) | (spread3(q[:, 1]) << 1) | spread3(q[:, 2])
    d = rays[:, 4:7]
    octant = ((d[:, 0] < 0).long() << 2) | ((d[:, 1] < 0).long() << 1) | (d[:, 2] < 0).long()
    if mode == "oct_origin":
        return (octant << 30) | mort
    if mode == "origin_oct":
        return (mort << 3) | octant
    if mode == "origin15_dir":   # coarse origin (5 bits/axis), then direction quantised 4 bits/axis
        dq = ((d / d.norm(dim=1, keepdim=True) * 0.5 + 0.5).clamp(0, 0.9999) * 16).long()
        dm = (spread3(dq[:, 0]) << 2) | (spread3(dq[:, 1]) << 1) | spread3(dq[:, 2])
        return ((mort >> 15) << 12) | dm
    raise ValueError(mode)


for name in sys.argv[1:] or ["soup1m", "flake10m"]:
    wl = WORKLOADS[name]
    tris = g.gen_scene(*wl.scene)
    with g.Bvh.build(tris) as bvh:
        box = bvh.bounds
        prim = g.gen_primary_rays(wl.camera(), wl.width, wl.height)
        _, hits = time_trace(bvh, prim, 0, 1)
        radius = 0.05 * float(np.linalg.norm(box[3:] - box[:3]))
        sets = {"ao_any": (g.gen_ao_rays(tris, prim, hits, 4, 2, radius), 1),
                "random4m": (g.gen_random_rays(box, 4, 0, 1 << 22), 0),
                "random32m": (g.gen_random_rays(box, 4, 0, 1 << 25), 0)}
        for kind, (rays, flags) in sets.items():
            ms0, ref = time_trace(bvh, rays, flags, 3)
            print(json.dumps(dict(scene=name, kind=kind, mode="as given", rays=rays.shape[0], ms=round(ms0, 4))), flush=True)
            for mode in ("oct_origin", "origin_oct", "origin15_dir"):
                perm = torch.argsort(keys(rays, box, mode))
                sorted_rays = rays[perm].contiguous()
                ms, out = time_trace(bvh, sorted_rays, flags, 3)
                same = bool(torch.equal(out.view(torch.int32), ref[perm].view(torch.int32)))
                print(json.dumps(dict(scene=name, kind=kind, mode=mode, ms=round(ms, 4),
                                      speedup=round(ms0 / ms, 3), same=same)), flush=True)
                del perm, sorted_rays
    del tris
    torch.cuda.empty_cache()
