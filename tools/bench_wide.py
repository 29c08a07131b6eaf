"""BVH2 vs BVH4 traversal timings on the BASELINE.json scenes (development tool)."""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpu_raytracer_b200 as g  # noqa: E402
from tools.probe import scene, time_trace  # noqa: E402

import itertools
LEAF = [int(x) for x in os.environ.get("LEAF_MAX", "4").split(",")]
for name, leaf_max in itertools.product(sys.argv[1:] or ["soup1m", "flake10m"], LEAF):
    tris = scene(name)
    g.set_param("wide.leaf_max", leaf_max)
    with g.Bvh.build(tris, wide=True, timings=True) as bvh:
        with g.Bvh.build(tris, wide=True, timings=True) as b2:   # second build: warm pool
            print(json.dumps({"scene": name, "leaf_max": leaf_max, "build_warm": b2.build_times()}), flush=True)
        print(json.dumps({"scene": name, "build": bvh.build_times()}), flush=True)
        box = bvh.bounds
        w, h = (1920, 1080) if name == "soup1m" else (3840, 2160)
        cam = g.camera.frame_box(box, w, h)
        rays = g.gen_primary_rays(cam, w, h)
        radius = 0.05 * float(np.linalg.norm(box[3:] - box[:3]))
        g.set_param("trace.wide", 0)
        _, hits = time_trace(bvh, rays)
        ao = g.gen_ao_rays(tris, rays, hits, 4, 2, radius)
        rr = g.gen_random_rays(box, 4, 0, 1 << 24)
        for kind, batch, flags in (("primary", rays, 0), ("ao_any", ao, 1), ("random", rr, 0)):
            res = {}
            for wide in (0, 1):
                g.set_param("trace.wide", wide)
                ms, out = time_trace(bvh, batch, flags)
                res[wide] = (ms, out)
            same = bool(torch.equal(res[0][1].view(torch.int32), res[1][1].view(torch.int32)))
            g.set_param("trace.wide", 1)
            _, nn, nt = bvh.trace_stats(batch, flags)
            g.set_param("trace.wide", 0)
            _, nn2, nt2 = bvh.trace_stats(batch, flags)
            n = batch.shape[0]
            print(json.dumps({"scene": name, "leaf_max": leaf_max, "kind": kind, "rays": n, "bvh2_ms": res[0][0],
                              "bvh4_ms": res[1][0], "speedup": res[0][0] / res[1][0],
                              "bvh4_mrays_s": n / res[1][0] / 1e3, "same": same,
                              "bvh2_nodes_tris": [nn2 / n, nt2 / n], "bvh4_nodes_tris": [nn / n, nt / n]}),
                  flush=True)
        g.set_param("trace.wide", 1)
    del tris
    torch.cuda.empty_cache()
