"""Tree-stage A/B of build_walk_kernel's shapes (development tool): rt_set_param("build.walk",
THREADS * 16 + LPT), LPT consecutive leaves per thread.  One JSON line per (scene, shape)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpu_raytracer_b200 as g
from gpu_raytracer_b200.configs import WORKLOADS
from tools.bench_sort import peak
shapes = [(256, 1), (128, 1), (512, 1), (256, 2), (128, 2), (128, 4), (256, 4), (64, 4)]
if len(sys.argv) > 1:  # e.g. 256x1,128x1,512x1
    shapes = [tuple(int(v) for v in a.split("x")) for a in sys.argv[1].split(",")]
for name in (sys.argv[2].split(",") if len(sys.argv) > 2 else ("soup1m", "flake10m", "terrain100m")):
    wl = WORKLOADS[name]
    tris = g.gen_scene(*wl.scene)
    n = int(tris.shape[0])
    for kb in ((30, 63) if name != "soup1m" else (30,)):
        for t, l in shapes:
            g.set_param("build.walk", t * 16 + l if (t, l) != (256, 1) else 0)
            runs = []
            for i in range(9):
                with g.Bvh.build(tris, key_bits=kb, timings=True) as b:
                    if i >= 2:
                        runs.append(b.build_times())
            med = {k: sorted(r[k] for r in runs)[len(runs) // 2] for k in runs[0]}
            tree = med["topology"] + med["refit"]
            byts = (216 if kb == 30 else 220) * n
            print(json.dumps({"scene": name, "key_bits": kb, "threads": t, "leaves_per_thread": l, "tree_ms": tree,
                              "tree_frac": byts / tree / 1e6 / peak(), "total_ms": med["total"]}), flush=True)
    del tris
    g.trim()
g.set_param("build.walk", 0)
