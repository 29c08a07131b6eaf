"""Times the traversal kernel with the top of the per-thread stack in shared memory
(`trace.sh_stack` levels) against the all-local stack (development tool)."""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpu_raytracer_b200 as g  # noqa: E402
from gpu_raytracer_b200.configs import WORKLOADS  # noqa: E402
from tools.tune_trace import time_trace  # noqa: E402


def main():
    names = sys.argv[1:] or ["soup1m", "flake10m"]
    levels = [int(x) for x in os.environ.get("LEVELS", "0,8,16,24").split(",")]
    for name in names:
        wl = WORKLOADS[name]
        tris = g.gen_scene(*wl.scene)
        with g.Bvh.build(tris) as bvh:
            box = bvh.bounds
            sets = {"primary": (g.gen_primary_rays(wl.camera(), wl.width, wl.height), 0)}
            g.set_param("trace.sh_stack", 0)
            _, hits = time_trace(bvh, sets["primary"][0], 0, 1)
            radius = 0.05 * float(np.linalg.norm(box[3:] - box[:3]))
            sets["ao_any"] = (g.gen_ao_rays(tris, sets["primary"][0], hits, 4, 2, radius), 1)
            sets["random"] = (g.gen_random_rays(box, 4, 0, 1 << 22), 0)
            for kind, (rays, flags) in sets.items():
                ref, ms0 = None, None
                for k in levels:
                    g.set_param("trace.sh_stack", k)
                    ms, out = time_trace(bvh, rays, flags, 5)
                    if ref is None:
                        ref, ms0 = out.clone(), ms
                    same = bool(torch.equal(out.view(torch.int32), ref.view(torch.int32)))
                    print(json.dumps(dict(scene=name, kind=kind, sh_stack=k, rays=rays.shape[0], ms=round(ms, 4),
                                          mrays_s=round(rays.shape[0] / ms / 1e3, 1),
                                          speedup=round(ms0 / ms, 3), same=same)), flush=True)
        del tris
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
