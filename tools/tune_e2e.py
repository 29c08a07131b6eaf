"""e2e trace_rays timing with pinned host buffers: zero-copy modes x staging chunk."""
import json, os, sys, time
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpu_raytracer_b200 as g
from gpu_raytracer_b200.configs import WORKLOADS
for name in sys.argv[1:] or ["soup1m"]:
    wl = WORKLOADS[name]
    tris = g.gen_scene(*wl.scene)
    rays = g.gen_primary_rays(wl.camera(), wl.width, wl.height)
    h_rays = rays.cpu().pin_memory()
    h_hits = torch.empty((rays.shape[0], 4)).pin_memory()
    with g.Bvh.build(tris) as bvh:
        for zc in (0, 1, 2, 3):
            for chunk in ((1 << 17, 1 << 18, 1 << 19, 1 << 20) if zc in (0, 2) else (0,)):
                g.set_param("trace.zero_copy", zc)
                g.set_param("trace.stage_chunk", chunk)
                ts = []
                for _ in range(6):
                    torch.cuda.synchronize()
                    t0 = time.perf_counter()
                    bvh.trace(h_rays, out=h_hits)
                    ts.append(time.perf_counter() - t0)
                ms = min(ts[1:]) * 1e3
                print(json.dumps(dict(scene=name, zero_copy=zc, chunk=chunk, ms=ms,
                                      mrays_s=rays.shape[0] / ms / 1e3)), flush=True)
