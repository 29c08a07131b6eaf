import os, sys, time
import numpy as np, torch
sys.path.insert(0, "/root/repo")
import gpu_raytracer_b200 as g
from gpu_raytracer_b200.configs import WORKLOADS
WL = WORKLOADS["flake10m"]
tris = g.gen_scene(*WL.scene)
bvh = g.Bvh.build(tris)
rays = g.gen_primary_rays(WL.camera(), WL.width, WL.height)
n = WL.width * WL.height
radius = float(np.float32(0.05) * np.float32(np.linalg.norm(bvh.bounds[3:] - bvh.bounds[:3])))
h_rays = rays.cpu().pin_memory(); h_hits = torch.empty((n, 4)).pin_memory(); h_rgb = torch.empty((n, 3)).pin_memory()
d_hits = torch.empty((n, 4), device="cuda"); d_rgb = torch.empty((n, 3), device="cuda")
cases = {"host": (h_rays, h_hits, h_rgb), "device": (rays, d_hits, d_rgb), "rays_host": (h_rays, d_hits, d_rgb),
         "out_host": (rays, h_hits, h_rgb)}
for name in (sys.argv[1:] or ["host"]):
    r, hh, gg = cases[name]
    print("case", name, file=sys.stderr)
    os.environ.pop("RT_RENDER_TIMELINE", None)
    for i in range(4):
        if i == 3: os.environ["RT_RENDER_TIMELINE"] = "1"
        torch.cuda.synchronize()
        t0 = time.perf_counter(); bvh.render_ao(tris, r, WL.ao_spp, WL.ao_seed, radius, hits=hh, rgb=gg); print("ms", (time.perf_counter() - t0) * 1e3, file=sys.stderr)
