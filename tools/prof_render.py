"""rt_render_ao on flake10m, 4K + 4 AO: wall time per frame by where the buffers live and by band
size, so the upload, the download and the compute can be told apart.  One JSON line per case.

  python tools/prof_render.py > gpurun_out/render.jsonl
"""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import gpu_raytracer_b200 as g  # noqa: E402
from gpu_raytracer_b200.configs import WORKLOADS  # noqa: E402

WL = WORKLOADS["flake10m"]
W, H, SPP, SEED = WL.width, WL.height, WL.ao_spp, WL.ao_seed


def main():
    tris = g.gen_scene(*WL.scene)
    bvh = g.Bvh.build(tris)
    cam = WL.camera()
    rays = g.gen_primary_rays(cam, W, H)
    n = W * H
    radius = float(np.float32(0.05) * np.float32(np.linalg.norm(bvh.bounds[3:] - bvh.bounds[:3])))
    h_rays = rays.cpu().pin_memory()
    h_hits = torch.empty((n, 4)).pin_memory()
    h_rgb = torch.empty((n, 3)).pin_memory()
    d_hits = torch.empty((n, 4), device="cuda")
    d_rgb = torch.empty((n, 3), device="cuda")
    cases = {"all host": (h_rays, h_hits, h_rgb), "rays host, out device": (h_rays, d_hits, d_rgb),
             "rays device, out host": (rays, h_hits, h_rgb), "all device": (rays, d_hits, d_rgb)}
    for band, taper in ((20, 3), (20, 1), (20, 0), (19, 3), (19, 0)):
        g.set_param("render.band", band)
        g.set_param("render.taper", taper)
        for name, (r, hh, gg) in cases.items():
            ts = []
            for it in range(12):
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                bvh.render_ao(tris, r, SPP, SEED, radius, hits=hh, rgb=gg)
                ts.append(time.perf_counter() - t0)
            ts = sorted(ts[2:])
            print(json.dumps({"band_log2": band, "taper": taper, "case": name, "ms_median": ts[len(ts) // 2] * 1e3, "ms_min": ts[0] * 1e3,
                              "mrays_s": n * (1 + SPP) / ts[len(ts) // 2] / 1e6}), flush=True)
    # copies alone, both directions at once
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    d_r = torch.empty_like(rays)
    for both in (False, True):
        ts = []
        for it in range(8):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            with torch.cuda.stream(s1):
                d_r.copy_(h_rays, non_blocking=True)
            if both:
                with torch.cuda.stream(s2):
                    h_hits.copy_(d_hits, non_blocking=True)
                    h_rgb.copy_(d_rgb, non_blocking=True)
            torch.cuda.synchronize()
            ts.append(time.perf_counter() - t0)
        ts = sorted(ts[2:])
        print(json.dumps({"case": "copies only" + (" (up + down together)" if both else " (up)"), "ms_median": ts[len(ts) // 2] * 1e3}), flush=True)
    bvh.close()


if __name__ == "__main__":
    main()
