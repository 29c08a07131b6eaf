"""Does the ORDER of a coherent frame's rays matter?  flake10m, 4K primary rays (+ 4 AO per pixel
made from them): row-major (what rt_gen_primary_rays gives: a warp holds a 32x1 strip) against the
same rays permuted into w x h pixel tiles and into Morton order.  Device time per launch, best of 7.

  python tools/tune_tile.py > gpurun_out/tile.jsonl
"""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import gpu_raytracer_b200 as g  # noqa: E402
from gpu_raytracer_b200.configs import WORKLOADS  # noqa: E402

WL = WORKLOADS["flake10m"]
W, H = WL.width, WL.height


def tile_perm(tw, th):
    """pixel indices in tile order: tiles row-major over the frame, pixels row-major inside a tile"""
    y, x = torch.meshgrid(torch.arange(H), torch.arange(W), indexing="ij")
    key = ((y // th) * ((W + tw - 1) // tw) + (x // tw)) * (tw * th) + (y % th) * tw + (x % tw)
    return torch.argsort(key.reshape(-1), stable=True)


def morton_perm():
    y, x = torch.meshgrid(torch.arange(H), torch.arange(W), indexing="ij")

    def spread(v):
        v = v.to(torch.int64)
        out = torch.zeros_like(v)
        for b in range(12):
            out |= ((v >> b) & 1) << (2 * b)
        return out
    return torch.argsort((spread(x) | (spread(y) << 1)).reshape(-1), stable=True)


def timed(fn, reps=7):
    best = 1e9
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best


def main():
    tris = g.gen_scene(*WL.scene)
    bvh = g.Bvh.build(tris)
    rays = g.gen_primary_rays(WL.camera(), W, H)
    n = W * H
    radius = float(np.float32(WL.ao_radius_frac) * np.float32(np.linalg.norm(bvh.bounds[3:] - bvh.bounds[:3])))
    orders = {"row-major": None, "8x4": tile_perm(8, 4), "4x8": tile_perm(4, 8), "16x2": tile_perm(16, 2),
              "8x8": tile_perm(8, 8), "32x32 of 8x4": None, "morton": morton_perm()}
    # two-level: 8x4 tiles inside 32x32 blocks
    y, x = torch.meshgrid(torch.arange(H), torch.arange(W), indexing="ij")
    key = (((y // 32) * ((W + 31) // 32) + x // 32) * 1024 + (((y % 32) // 4) * 4 + (x % 32) // 8) * 32 + (y % 4) * 8 + x % 8)
    orders["32x32 of 8x4"] = torch.argsort(key.reshape(-1), stable=True)
    hits = torch.empty((n, 4), device="cuda")
    ao_hits = torch.empty((n * WL.ao_spp, 4), device="cuda")
    base = None
    for name, perm in orders.items():
        r = rays if perm is None else rays[perm.cuda()].contiguous()
        ms = timed(lambda: bvh.trace_async(r, hits))
        torch.cuda.synchronize()
        ao = g.gen_ao_rays(tris, r, hits, WL.ao_spp, WL.ao_seed, radius)
        ms_ao = timed(lambda: bvh.trace_async(ao, ao_hits, g.RT_TRACE_ANY))
        _, nodes, tests = bvh.trace_stats(r)
        if base is None:
            base = (ms, ms_ao)
        print(json.dumps({"order": name, "primary_ms": ms, "ao_ms": ms_ao, "primary_speedup": base[0] / ms,
                          "ao_speedup": base[1] / ms_ao, "mrays_s": n * (1 + WL.ao_spp) / (ms + ms_ao) / 1e3,
                          "nodes_per_ray": nodes / n}), flush=True)
    bvh.close()


if __name__ == "__main__":
    main()
