"""Streamed trace_rays under a profiler that replays kernels (development check): the kernel
sees no upload progress, gives up after 2 s, and the batch is traced again through staged
chunks.  usage: ncu --set full -k regex:trace_kernel_dyn python tools/stall_check.py"""
import os
import sys
import time

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpu_raytracer_b200 as g  # noqa: E402
from gpu_raytracer_b200.configs import WORKLOADS  # noqa: E402

wl = WORKLOADS["soup1m"]
tris = g.gen_scene(*wl.scene)
rays = g.gen_primary_rays(wl.camera(), wl.width, wl.height)
h_rays = rays.cpu().pin_memory()
h_hits = torch.empty((rays.shape[0], 4)).pin_memory()
with g.Bvh.build(tris) as bvh:
    want = bvh.trace(rays).cpu()
    for it in range(2):
        t0 = time.perf_counter()
        bvh.trace(h_rays, out=h_hits)
        print("call", it, "ms", round((time.perf_counter() - t0) * 1e3, 2), "same",
              bool(torch.equal(h_hits.view(torch.int32), want.view(torch.int32))), flush=True)
