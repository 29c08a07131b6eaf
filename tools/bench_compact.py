"""The 64-byte binary nodes against the 32-byte compact ones (RT_BUILD_COMPACT) on the bench's
trace launches: flake10m primary / AO / random rays (plain and reordered), soup1m primary and
random, with the cost of packing.  Every pair of runs is compared bit for bit.  One JSON line
per measurement.
usage: bench_compact.py [reps]"""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpu_raytracer_b200 as g  # noqa: E402
from gpu_raytracer_b200.configs import WORKLOADS  # noqa: E402

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 5
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)


def timed(bvh, rays, hits, flags=0):
    ms = []
    for _ in range(reps + 1):
        flush.zero_()
        e0.record()
        bvh.trace_async(rays, hits, flags)
        e1.record()
        torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    ms = sorted(ms[1:])
    return ms[len(ms) // 2]


def run(name):
    wl = WORKLOADS[name]
    tris = g.gen_scene(*wl.scene)
    n = int(tris.shape[0])
    t_build = {}
    for compact in (False, True):
        ts = []
        for _ in range(5):
            torch.cuda.synchronize()
            e0.record()
            b = g.Bvh.build(tris, compact=compact)
            e1.record()
            torch.cuda.synchronize()
            ts.append(e0.elapsed_time(e1))
            b.close()
        t_build[compact] = min(ts[1:])
    print(json.dumps({"scene": name, "n_tris": n, "build_ms": t_build[False], "build_compact_ms": t_build[True],
                      "pack_ms": t_build[True] - t_build[False]}), flush=True)
    with g.Bvh.build(tris, compact=True) as bvh:
        rays = g.gen_primary_rays(wl.camera(), wl.width, wl.height)
        hits = torch.empty((rays.shape[0], 4), dtype=torch.float32, device="cuda")
        box = bvh.bounds
        sets = [("primary", rays, hits, 0)]
        g.set_param("trace.compact", 0)
        bvh.trace_async(rays, hits)
        if wl.ao_spp:
            radius = float(np.float32(wl.ao_radius_frac) * np.float32(np.linalg.norm(box[3:] - box[:3])))
            ao = g.gen_ao_rays(tris, rays, hits, wl.ao_spp, wl.ao_seed, radius)
            sets.append(("ao", ao, torch.empty((ao.shape[0], 4), dtype=torch.float32, device="cuda"), g.RT_TRACE_ANY))
        rr = g.gen_random_rays(box, 4, 0, 1 << 24)
        rh = torch.empty((rr.shape[0], 4), dtype=torch.float32, device="cuda")
        sets.append(("random", rr, rh, 0))
        sets.append(("random_reorder", rr, rh, g.RT_TRACE_REORDER))
        for label, r, h, flags in sets:
            out = {"scene": name, "rays": label, "n_rays": int(r.shape[0])}
            ref = None
            for compact in (0, 1):
                g.set_param("trace.compact", compact)
                ms = timed(bvh, r, h, flags)
                key = "compact" if compact else "bvh2"
                out[key + "_ms"] = ms
                out[key + "_mrays_s"] = r.shape[0] / ms / 1e3
                if ref is None:
                    ref = h.clone()
                else:
                    out["hits_equal"] = bool(torch.equal(ref.view(torch.int32), h.view(torch.int32)))
                if not (flags & g.RT_TRACE_REORDER):
                    _, nd, nt = bvh.trace_stats(r, flags & g.RT_TRACE_ANY)
                    out[key + "_nodes_per_ray"] = nd / r.shape[0]
                    out[key + "_tris_per_ray"] = nt / r.shape[0]
            out["speedup"] = out["bvh2_ms"] / out["compact_ms"]
            print(json.dumps(out), flush=True)
    g.set_param("trace.compact", 1)


if __name__ == "__main__":
    for name in ("flake10m", "soup1m"):
        run(name)
