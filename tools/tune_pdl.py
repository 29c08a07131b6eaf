"""A/B of programmatic dependent launch on the build chain (`build.pdl`): events around
build_bvh on device-resident triangles, L2 flushed before every build (development tool)."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpu_raytracer_b200 as g  # noqa: E402
from gpu_raytracer_b200.configs import WORKLOADS  # noqa: E402


def time_build(tris, key_bits, reps):
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ts = []
    for _ in range(reps):
        flush.zero_()
        e0.record()
        bvh = g.Bvh.build(tris, key_bits=key_bits)
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
        bvh.close()
    ts.sort()
    return ts[0], ts[len(ts) // 2]


def main():
    names = sys.argv[1:] or ["cornell", "soup1m", "flake10m"]
    for name in names:
        wl = WORKLOADS[name]
        tris = g.gen_scene(*wl.scene)
        for rnd in range(2):
            for pdl in (0, 1):
                g.set_param("build.pdl", pdl)
                time_build(tris, wl.key_bits, 3)
                best, med = time_build(tris, wl.key_bits, 30)
                print(json.dumps(dict(scene=name, n=int(tris.shape[0]), pdl=pdl, round=rnd,
                                      best_ms=round(best, 4), median_ms=round(med, 4))), flush=True)
        del tris
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
