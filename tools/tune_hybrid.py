"""63-bit build with and without the hybrid sort order (`sort.hybrid`): stage times and
equality of the sorted arrays (development tool)."""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpu_raytracer_b200 as g  # noqa: E402
from tools.probe import scene  # noqa: E402

for name in sys.argv[1:] or ["terrain100m", "flake10m", "soup1m"]:
    tris = scene(name)
    ref = None
    for hyb in (0, 1, 0, 1):
        g.set_param("sort.hybrid", hyb)
        best = None
        for _ in range(3):
            with g.Bvh.build(tris, key_bits=63, timings=True) as b:
                t = b.build_times()
            if best is None or t["total"] < best["total"]:
                best = t
        same = None
        if name != "terrain100m" or ref is None or hyb == 1:
            with g.Bvh.build(tris, key_bits=63, keep_keys=True) as b:
                keys, idx = b.read_sorted()
            if ref is None:
                ref = (keys, idx)
            else:
                same = bool(np.array_equal(keys, ref[0]) and np.array_equal(idx, ref[1]))
            del keys, idx
        print(json.dumps(dict(scene=name, hybrid=hyb, sort_ms=round(best["sort"], 4),
                              tree_ms=round(best["topology"] + best["refit"], 4),
                              total_ms=round(best["total"], 4), same_as_first=same)), flush=True)
    g.set_param("sort.hybrid", 1)
    del tris, ref
    torch.cuda.empty_cache()
