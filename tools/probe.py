"""Quick device-side timing probe (development tool, not the bench contract).

Prints per-stage build times, traversal rates and sort-variant timings for the
BASELINE.json configs so one gpurun call answers "where does the time go"."""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpu_raytracer_b200 as g  # noqa: E402

OUT = []


def emit(**kw):
    OUT.append(kw)
    print(json.dumps(kw), flush=True)


def time_build(tris, key_bits, reps=5):
    best = None
    for _ in range(reps):
        with g.Bvh.build(tris, key_bits=key_bits, timings=True) as b:
            t = b.build_times()
        if best is None or t["total"] < best["total"]:
            best = t
    return best


def time_trace(bvh, rays, flags=0, reps=5):
    out = torch.empty((rays.shape[0], 4), dtype=torch.float32, device="cuda")
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    best = 1e30
    for _ in range(reps):
        e0.record()
        bvh.trace_async(rays, out, flags)
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best, out


def scene(name):
    if name == "soup1m":
        return g.gen_scene(g.RT_SCENE_SOUP, 1000000, 0, 1)
    if name == "flake10m":
        return g.gen_scene(g.RT_SCENE_SPHEREFLAKE, 4, 13, 2)
    if name == "terrain100m":
        return g.gen_scene(g.RT_SCENE_TERRAIN, 7072, 0, 3)
    raise ValueError(name)


def main():
    which = sys.argv[1:] or ["soup1m", "flake10m", "sort"]
    for name in which:
        if name == "sort":
            for key_bits in (30, 63):
                for n in (1000000, 10000000):
                    for bits, variant in [(8, 0), (10, 0), (10, 1), (11, 0), (11, 1), (9, 0), (9, 1), (8, 1)]:
                        g.set_param("sort.bits", bits)
                        g.set_param("sort.variant", variant)
                        ts = []
                        for rep in range(4):
                            gen = torch.Generator(device="cuda").manual_seed(rep)
                            if key_bits == 30:
                                k = torch.randint(0, 1 << 30, (n,), device="cuda", dtype=torch.int32, generator=gen)
                            else:
                                k = torch.randint(0, 1 << 62, (n,), device="cuda", dtype=torch.int64, generator=gen)
                            v = torch.arange(n, device="cuda", dtype=torch.int32)
                            ts.append(g.sort_pairs(k, v, key_bits))
                        ms = min(ts)
                        alg = n * (68 if key_bits == 30 else 200)  # SURVEY 8d figure for A3
                        emit(probe="sort", key_bits=key_bits, n=n, bits=bits, variant=variant, ms=ms,
                             gpairs_s=n / ms / 1e6, alg_gbs=alg / ms / 1e6)
            g.set_param("sort.variant", 0)
            g.set_param("sort.bits", 0)
            continue
        if name.startswith("sortscene:"):
            sc = name.split(":")[1]
            tris = scene(sc)
            for key_bits in (30, 63):
                for bits, variant in [(8, 0), (8, 1), (9, 0), (9, 1), (10, 0), (10, 1), (11, 0), (11, 1)]:
                    g.set_param("sort.bits", bits)
                    g.set_param("sort.variant", variant)
                    t = time_build(tris, key_bits, reps=4)
                    emit(probe="sortscene", scene=sc, key_bits=key_bits, bits=bits, variant=variant,
                         sort_ms=t["sort"], morton_ms=t["morton"], total_ms=t["total"])
            g.set_param("sort.variant", 0)
            g.set_param("sort.bits", 0)
            del tris
            continue
        t0 = time.time()
        tris = scene(name)
        torch.cuda.synchronize()
        n = tris.shape[0]
        emit(probe="gen", scene=name, n=n, s=time.time() - t0)
        for key_bits in ((63,) if name == "terrain100m" else (30, 63)):
            t = time_build(tris, key_bits, reps=3 if n > 2e7 else 5)
            emit(probe="build", scene=name, key_bits=key_bits, n=n, mtri_s=n / t["total"] / 1e3, **t)
        if name == "terrain100m":
            continue
        with g.Bvh.build(tris) as bvh:
            box = bvh.bounds
            w, h = (1920, 1080) if name == "soup1m" else (3840, 2160)
            cam = g.camera.frame_box(box, w, h)
            rays = g.gen_primary_rays(cam, w, h)
            for bps in (4, 6, 8):
                g.set_param("trace.blocks_per_sm", bps)
                ms, hits = time_trace(bvh, rays)
                emit(probe="trace", scene=name, kind="primary", rays=rays.shape[0], bps=bps, ms=ms,
                     mrays_s=rays.shape[0] / ms / 1e3,
                     hit_frac=float((g.hit_prims(hits) >= 0).float().mean()))
            g.set_param("trace.blocks_per_sm", 0)
            radius = 0.05 * float(np.linalg.norm(box[3:] - box[:3]))
            ao = g.gen_ao_rays(tris, rays, hits, 4, 2, radius)
            ms, occ = time_trace(bvh, ao, 1)
            emit(probe="trace", scene=name, kind="ao_any", rays=ao.shape[0], ms=ms,
                 mrays_s=ao.shape[0] / ms / 1e3,
                 occluded=float((g.hit_prims(occ) == g.RT_ANY_HIT_PRIM).float().mean()))
            del ao, occ
            rr = g.gen_random_rays(box, 4, 0, 1 << 24)
            ms, hits = time_trace(bvh, rr)
            emit(probe="trace", scene=name, kind="random", rays=rr.shape[0], ms=ms,
                 mrays_s=rr.shape[0] / ms / 1e3,
                 hit_frac=float((g.hit_prims(hits) >= 0).float().mean()))
        del tris
        torch.cuda.empty_cache()
    os.makedirs("gpurun_out", exist_ok=True)
    with open("gpurun_out/probe.json", "w") as f:
        json.dump(OUT, f, indent=1)


if __name__ == "__main__":
    main()
