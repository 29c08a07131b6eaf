"""The trace launches of bench.py's default workload, once each (the command ncu wraps):
flake10m with the bench camera: 3840x2160 primary rays (closest hit), their 4 AO rays per pixel
(any-hit), and a 2^24-ray window of config 5's random rays, plain and with RT_TRACE_REORDER.
usage: prof_bench_trace.py [wide8 | compact]   (trace the compressed 8-wide layout / the 32-byte
binary nodes instead of the 64-byte BVH2)"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpu_raytracer_b200 as g  # noqa: E402
from gpu_raytracer_b200.configs import WORKLOADS  # noqa: E402

wl = WORKLOADS["flake10m"]
wide8 = "wide8" in sys.argv[1:]
tris = g.gen_scene(*wl.scene)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
kw = dict(wide8=True) if wide8 else (dict(compact=True) if "compact" in sys.argv[1:] else {})
with g.Bvh.build(tris, **kw) as bvh:
    cam = wl.camera()
    rays = g.gen_primary_rays(cam, wl.width, wl.height)
    hits = torch.empty((rays.shape[0], 4), dtype=torch.float32, device="cuda")
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)

    def timed(name, r, h, flags=0):
        flush.zero_()
        e0.record()
        bvh.trace_async(r, h, flags)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        print(name, f"{ms:.3f} ms", f"{r.shape[0] / ms / 1e3:.1f} Mrays/s", flush=True)

    timed("primary", rays, hits)
    box = bvh.bounds
    radius = wl.ao_radius_frac * float(np.linalg.norm(box[3:] - box[:3]))
    ao = g.gen_ao_rays(tris, rays, hits, wl.ao_spp, wl.ao_seed, radius)
    ao_hits = torch.empty((ao.shape[0], 4), dtype=torch.float32, device="cuda")
    timed("ao", ao, ao_hits, g.RT_TRACE_ANY)
    rr = g.gen_random_rays(box, 4, 0, 1 << 24)
    rh = torch.empty((rr.shape[0], 4), dtype=torch.float32, device="cuda")
    timed("random", rr, rh)
    timed("random_reorder", rr, rh, g.RT_TRACE_REORDER)
torch.cuda.synchronize()
