"""Digit width / CTA shape sweep of the radix sort on 63-bit keys (development tool):
rt_sort_pairs on random 63-bit keys, and the sort stage of a full build."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpu_raytracer_b200 as g  # noqa: E402
from tools.probe import scene  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "terrain100m"
key_bits = int(sys.argv[2]) if len(sys.argv) > 2 else 63
tris = scene(name)
for variant in (0, 1):
    for bits in ((9, 10, 11) if key_bits == 63 else (8, 10)):
        g.set_param("sort.variant", variant)
        g.set_param("sort.bits", bits)
        best = None
        for _ in range(3):
            with g.Bvh.build(tris, key_bits=key_bits, timings=True) as b:
                t = b.build_times()
            if best is None or t["sort"] < best["sort"]:
                best = t
        print(json.dumps(dict(scene=name, key_bits=key_bits, variant=variant, bits=bits,
                              sort_ms=round(best["sort"], 4), morton_ms=round(best["morton"], 4),
                              total_ms=round(best["total"], 4))), flush=True)
g.set_param("sort.variant", 0)
g.set_param("sort.bits", 0)
