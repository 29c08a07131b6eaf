"""Sort-stage A/B of the scatter kernel's variants (development tool): sort stage of build_bvh
(30-bit keys) and rt_sort_pairs, per `sort.variant`.  One JSON line per measurement.
variants: 0 = 512 threads x 15 keys (default), 1 = 256 x 16, 2 = 512 x 9, 3 = 512 x 9 with
TMA-staged record tiles (cp.async.bulk + mbarrier)."""
import json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpu_raytracer_b200 as g
from gpu_raytracer_b200.configs import WORKLOADS
from tools.bench_sort import keys_for, peak

variants = [int(v) for v in sys.argv[1:]] or [0, 1, 2, 3]
for name in ("soup1m", "flake10m", "terrain100m"):
    tris = g.gen_scene(*WORKLOADS[name].scene)
    n = int(tris.shape[0])
    for v in variants:
        g.set_param("sort.variant", v)
        runs = []
        for i in range(9):
            with g.Bvh.build(tris, key_bits=30, timings=True) as b:
                if i >= 2:
                    runs.append(b.build_times())
        med = {k: sorted(r[k] for r in runs)[len(runs) // 2] for k in runs[0]}
        print(json.dumps({"scene": name, "n": n, "variant": v, "sort_ms": med["sort"], "total_ms": med["total"],
                          "sort_frac": 68.0 * n / med["sort"] / 1e6 / peak()}), flush=True)
    del tris
    g.trim()
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for n in (1_000_000, 9_979_112, 100_026_368):
    for v in variants:
        g.set_param("sort.variant", v)
        ms = []
        for it in range(9):
            k = keys_for(n, 30); val = torch.arange(n, dtype=torch.int32, device="cuda")
            flush.fill_(it)
            t = g.sort_pairs(k, val, 30)
            if it >= 2: ms.append(t)
        ok = bool((k[1:] >= k[:-1]).all())
        ms.sort()
        print(json.dumps({"rt_sort_pairs": n, "variant": v, "ms": ms[len(ms) // 2], "sorted": ok}), flush=True)
g.set_param("sort.variant", 0)
