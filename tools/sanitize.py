"""A small end-to-end pass over every kernel for compute-sanitizer (development tool):
  compute-sanitizer --tool memcheck python tools/sanitize.py
Scenes are small because the sanitizer slows kernels down by 10-100x."""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpu_raytracer_b200 as g  # noqa: E402

for kind, p0, p1, seed, kb in [(g.RT_SCENE_CORNELL, 0, 0, 0, 30), (g.RT_SCENE_SOUP, 30001, 0, 1, 30),
                               (g.RT_SCENE_SPHEREFLAKE, 2, 4, 0, 63), (g.RT_SCENE_TERRAIN, 64, 16, 3, 63)]:
    tris = g.gen_scene(kind, p0, p1, seed)
    with g.Bvh.build(tris, key_bits=kb, keep_keys=True, timings=True, wide=True) as b:
        box = b.bounds
        rays = g.gen_random_rays(box, 4, 0, 40000)
        cam = g.camera.frame_box(box, 96, 64)
        prim = g.gen_primary_rays(cam, 96, 64)
        for flags in (0, 1, 2, 3):
            h = b.trace(rays, flags)
            g.set_param("trace.wide", 0)
            h2 = b.trace(rays, flags)
            g.set_param("trace.variant", 0)
            h3 = b.trace(rays, flags)
            g.set_param("trace.variant", 1)
            g.set_param("trace.wide", 1)
            assert torch.equal(h.view(torch.int32), h2.view(torch.int32))
            assert torch.equal(h.view(torch.int32), h3.view(torch.int32))
        hits = b.trace(prim)
        ao = g.gen_ao_rays(tris, prim, hits, 4, 2, 0.05 * float(np.linalg.norm(box[3:] - box[:3])))
        occ = b.trace(ao, 1)
        g.shade(tris, prim, hits, occ, 4)
        b.trace_stats(rays)
        b.read_sorted(); b.read_topology(); b.read_boxes(); b.read_wide()
        b.refit(tris)
        b.trace(rays.cpu().numpy())           # staged host path
    # round 2: compressed 8-wide layout, reorder, camera entry, sharded scene + merge, no-alloc sort
    with g.Bvh.build(tris, key_bits=kb, wide8=True) as b8:
        for flags in (0, 1, 2, g.RT_TRACE_REORDER):
            b8.trace(g.gen_random_rays(box, 4, 0, 70000), flags)
        b8.read_wide8()
        b8.trace_camera(cam, 96, 64)
        b8.refit(tris)
    host = tris.cpu().numpy()
    with g.ShardedBvh.build(host, [0, 0, 0]) as sb:
        sb.trace(rays.cpu().numpy())
    with g.MultiBvh.build(host, [0, 0], strategy=g.RT_MULTI_BROADCAST) as mb:
        mb.trace(rays.cpu().numpy())
        mb.trace_random(box, 4, 0, 50000, g.RT_TRACE_REORDER, batch=20000)
    print("ok", kind, int(tris.shape[0]), flush=True)
k = torch.randint(0, 1 << 30, (100003,), device="cuda", dtype=torch.int32)
g.sort_pairs(k, torch.arange(100003, device="cuda", dtype=torch.int32), 30)
torch.cuda.synchronize()
print("sanitize pass done")

# round 2: the sort without allocations, twice on the shared static scratch; a BVH file round trip
for n in (100003, 7001):
    k2 = torch.randint(0, 1 << 30, (n,), device="cuda", dtype=torch.int32)
    v2 = torch.arange(n, device="cuda", dtype=torch.int32)
    g.sort_pairs_static(k2, v2)
    torch.cuda.synchronize()
    assert bool((k2[1:] >= k2[:-1]).all())
import tempfile
with tempfile.TemporaryDirectory() as d:
    t = g.gen_scene(g.RT_SCENE_SOUP, 5000, 0, 3)
    with g.Bvh.build(t) as b:
        b.save(os.path.join(d, "a.bvh"))
    with g.Bvh.load(os.path.join(d, "a.bvh")) as b:
        b.widen8()
        b.trace(g.gen_random_rays(b.bounds, 4, 0, 20000))
print("round-2 paths ok", flush=True)
