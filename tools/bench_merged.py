"""Sharded BUILD (build_bvh_sharded / trace_rays_merged): the scene cut over the visible GPUs, every
part traces the whole batch, the hits merged by (t, prim).  Times the host-to-host call with the
exchange and the merge fused (every device merges its slice out of peer memory and downloads it)
against the gather on the first device.  One JSON line per variant.
usage: bench_merged.py [n_rays_log2]"""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpu_raytracer_b200 as g  # noqa: E402
from gpu_raytracer_b200.configs import WORKLOADS  # noqa: E402

wl = WORKLOADS["flake10m"]
n_dev = torch.cuda.device_count()
n_rays = 1 << (int(sys.argv[1]) if len(sys.argv) > 1 else 24)
tris = g.gen_scene(*wl.scene).cpu().numpy()
v = tris.reshape(-1, 4)[:, :3]
box = np.concatenate([v.min(0), v.max(0)]).astype(np.float32)
rays = torch.from_numpy(g.gen_random_rays(box, 4, 0, n_rays).cpu().numpy()).pin_memory()
hits = torch.empty((n_rays, 4), dtype=torch.float32).pin_memory()
with g.ShardedBvh.build(tris, list(range(n_dev))) as sb:
    ref = None
    for fused in (1, 0, 1, 0):
        g.set_param("merge.fused", fused)
        ts = []
        for _ in range(4):
            t0 = time.perf_counter()
            sb.trace(rays, out=hits)
            ts.append(time.perf_counter() - t0)
        if ref is None:
            ref = hits.clone()
        print(json.dumps({"n_gpus": n_dev, "n_rays": n_rays, "fused": fused, "ms": min(ts[1:]) * 1e3,
                          "mrays_s": n_rays / min(ts[1:]) / 1e6,
                          "equal": bool(torch.equal(ref.view(torch.int32), hits.view(torch.int32)))}), flush=True)
# build_bvh_multi, redundant strategy, pinned host triangles: every device uploads the whole
# array (0) against 1/n each + an all-gather over peer memory (1)
h_tris = torch.from_numpy(tris).pin_memory()
for fused in (1, 0, 1, 0):
    g.set_param("merge.fused", fused)
    ts = []
    for _ in range(4):
        t0 = time.perf_counter()
        mb = g.MultiBvh.build(h_tris, list(range(n_dev)))
        ts.append(time.perf_counter() - t0)
        mb.close()
    print(json.dumps({"n_gpus": n_dev, "call": "build_bvh_multi (redundant, host triangles)", "n_tris": int(tris.shape[0]),
                      "fused": fused, "ms": min(ts[1:]) * 1e3}), flush=True)
g.set_param("merge.fused", 1)
