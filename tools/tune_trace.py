"""Sweeps the traversal kernel's knobs on the bench workloads (development tool)."""
import itertools
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpu_raytracer_b200 as g  # noqa: E402
from gpu_raytracer_b200.configs import WORKLOADS  # noqa: E402


def time_trace(bvh, rays, flags, reps=3):
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


def main():
    names = sys.argv[1:] or ["soup1m", "flake10m"]
    full = os.environ.get("SWEEP", "1") == "1"
    for name in names:
        wl = WORKLOADS[name]
        tris = g.gen_scene(*wl.scene)
        with g.Bvh.build(tris) as bvh:
            box = bvh.bounds
            sets = {"primary": (g.gen_primary_rays(wl.camera(), wl.width, wl.height), 0)}
            g.set_param("trace.variant", 0)
            _, hits = time_trace(bvh, sets["primary"][0], 0, 1)
            radius = 0.05 * float(np.linalg.norm(box[3:] - box[:3]))
            sets["ao_any"] = (g.gen_ao_rays(tris, sets["primary"][0], hits, 4, 2, radius), 1)
            sets["random"] = (g.gen_random_rays(box, 4, 0, 1 << 22), 0)
            for kind, (rays, flags) in sets.items():
                g.set_param("trace.variant", 0)
                ms0, ref = time_trace(bvh, rays, flags)
                print(json.dumps(dict(scene=name, kind=kind, variant=0, rays=rays.shape[0], ms=ms0,
                                      mrays_s=rays.shape[0] / ms0 / 1e3,
                                      hit=float((g.hit_prims(ref) >= 0).float().mean()))), flush=True)
                g.set_param("trace.variant", 1)
                grid = (itertools.product((64, 128, 256), (2, 4, 8, 16), (4, 8, 12, 16, 24)) if full
                        else [(128, 4, 8)])
                res = []
                for chunk, refill, leaf in grid:
                    g.set_param("trace.chunk", chunk)
                    g.set_param("trace.refill_min", refill)
                    g.set_param("trace.leaf_min", leaf)
                    ms, out = time_trace(bvh, rays, flags, 2)
                    same = bool(torch.equal(out.view(torch.int32), ref.view(torch.int32)))
                    res.append((ms, chunk, refill, leaf, same))
                res.sort()
                for ms, chunk, refill, leaf, same in res[:6] + res[-2:]:
                    print(json.dumps(dict(scene=name, kind=kind, variant=1, chunk=chunk, refill=refill,
                                          leaf=leaf, ms=ms, mrays_s=rays.shape[0] / ms / 1e3,
                                          speedup=ms0 / ms, same=same)), flush=True)
                assert all(r[4] for r in res), "variant 1 differs from variant 0"
        del tris
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
