"""e2e trace_rays timing with pinned host buffers: streamed single-launch mode vs staged
chunks vs zero-copy (development tool).  usage: tune_e2e.py [scene ...]"""
import json
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpu_raytracer_b200 as g  # noqa: E402
from gpu_raytracer_b200.configs import WORKLOADS  # noqa: E402

for name in sys.argv[1:] or ["soup1m"]:
    wl = WORKLOADS[name]
    tris = g.gen_scene(*wl.scene)
    rays = g.gen_primary_rays(wl.camera(), wl.width, wl.height)
    h_rays = rays.cpu().pin_memory()
    h_hits = torch.empty((rays.shape[0], 4)).pin_memory()
    with g.Bvh.build(tris) as bvh:
        want = bvh.trace(rays).cpu()
        modes = [("streamed", 1, 0, 0)] + [("staged", 0, 0, c) for c in (1 << 18, 1 << 19, 1 << 20)] + \
                [("zero-copy rays", 0, 1, 0), ("zero-copy hits", 0, 2, 1 << 20)]
        if os.environ.get("RT_ONLY_STREAMED"):
            modes = modes[:1]
        for label, streamed, zc, chunk in modes:
            g.set_param("trace.streamed", streamed)
            g.set_param("trace.zero_copy", zc)
            g.set_param("trace.stage_chunk", chunk)
            ts = []
            h_hits.zero_()  # once: freshly zeroed (cache-dirty) host lines slow the next download
            for _ in range(8):
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                bvh.trace(h_rays, out=h_hits)
                ts.append(time.perf_counter() - t0)
            same = bool(torch.equal(h_hits.view(torch.int32), want.view(torch.int32)))
            ms = min(ts[1:]) * 1e3
            print(json.dumps(dict(scene=name, mode=label, chunk=chunk, ms=round(ms, 3),
                                  mrays_s=round(rays.shape[0] / ms / 1e3, 1), same=same)), flush=True)
        g.set_param("trace.streamed", 1)
        g.set_param("trace.zero_copy", 0)
        g.set_param("trace.stage_chunk", 0)
