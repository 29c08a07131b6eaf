"""Development tool: the 8-wide layout on flake10m primary rays: whole frame, quarters of it and
knob variations (the command ncu wraps with -k regex:trace_wide8)."""
import os, sys, json
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpu_raytracer_b200 as g
from gpu_raytracer_b200.configs import WORKLOADS
wl = WORKLOADS["flake10m"]
tris = g.gen_scene(*wl.scene)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
with g.Bvh.build(tris, wide8=True) as bvh:
    rays = g.gen_primary_rays(wl.camera(), wl.width, wl.height)
    hits = torch.empty((rays.shape[0], 4), dtype=torch.float32, device="cuda")
    def t(r, h, reps=3):
        best = 1e9
        for _ in range(reps):
            e0.record(); bvh.trace_async(r, h); e1.record(); torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        return best
    print("full", t(rays, hits), flush=True)
    if "quick" in sys.argv:
        sys.exit(0)
    n = rays.shape[0]
    for q in range(4):
        a, b = q * n // 4, (q + 1) * n // 4
        print("quarter", q, t(rays[a:b], hits[a:b]), flush=True)
    hp = g.hit_prims(hits)
    hit_rays = rays[hp >= 0].contiguous(); miss_rays = rays[hp < 0].contiguous()
    print("hits only", hit_rays.shape[0], t(hit_rays, hits[:hit_rays.shape[0]]), flush=True)
    print("misses only", miss_rays.shape[0], t(miss_rays, hits[:miss_rays.shape[0]]), flush=True)
    perm = torch.randperm(n, device="cuda")
    print("shuffled", t(rays[perm].contiguous(), hits), flush=True)
    for name, vals in (("trace.leaf_min", (1, 4, 16, 32)), ("trace.refill_min", (1, 16, 32)), ("trace.chunk", (32, 256))):
        for v in vals:
            g.set_param(name, v)
            print(name, v, t(rays, hits), flush=True)
        g.set_param(name, 0)
    g.set_param("trace.wide", 0)
    print("bvh2 full", t(rays, hits), "misses", t(miss_rays, hits[:miss_rays.shape[0]]), flush=True)
