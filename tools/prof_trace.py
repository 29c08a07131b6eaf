"""Traces the bench workloads a few times (development tool: the command ncu wraps).
usage: prof_trace.py <scene> [wide] [reps]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpu_raytracer_b200 as g  # noqa: E402
from tools.probe import scene, time_trace  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "soup1m"
wide = len(sys.argv) > 2 and sys.argv[2] == "wide"
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 2
tris = scene(name)
with g.Bvh.build(tris, wide=wide) as bvh:
    box = bvh.bounds
    w, h = (1920, 1080) if name == "soup1m" else (3840, 2160)
    cam = g.camera.frame_box(box, w, h)
    rays = g.gen_primary_rays(cam, w, h)
    ms, hits = time_trace(bvh, rays, reps=reps)
    print("primary", ms, rays.shape[0] / ms / 1e3, flush=True)
    radius = 0.05 * float(np.linalg.norm(box[3:] - box[:3]))
    ao = g.gen_ao_rays(tris, rays, hits, 4, 2, radius)
    ms, _ = time_trace(bvh, ao, 1, reps=reps)
    print("ao", ms, ao.shape[0] / ms / 1e3, flush=True)
    rr = g.gen_random_rays(box, 4, 0, 1 << 23)
    ms, _ = time_trace(bvh, rr, reps=reps)
    print("random", ms, rr.shape[0] / ms / 1e3, flush=True)
torch.cuda.synchronize()
