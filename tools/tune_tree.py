"""A/B of the tree-build kernels (`build.tree_variant`: 0 = walk with meetings, 1 / 2 = nodes
that compute themselves, 128 / 256 leaves per CTA): canonical arrays must be identical, then
events around build_bvh (development tool)."""
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpu_raytracer_b200 as g  # noqa: E402
from gpu_raytracer_b200.configs import WORKLOADS  # noqa: E402
from tools.tune_pdl import time_build  # noqa: E402

L = g._lib if hasattr(g, "_lib") else None


def arrays(tris, key_bits):
    with g.Bvh.build(tris, key_bits=key_bits, keep_keys=True) as b:
        out = list(b.read_topology()) + [b.read_boxes().view(np.uint32)]
        out.append(np.asarray(b.bounds).view(np.uint32))
    return out


def check(tris, key_bits, variants):
    g.set_param("build.tree_variant", 0)
    ref = arrays(tris, key_bits)
    ok = {}
    for v in variants:
        g.set_param("build.tree_variant", v)
        got = arrays(tris, key_bits)
        ok[v] = all(np.array_equal(a, b) for a, b in zip(ref, got))
    g.set_param("build.tree_variant", 0)
    return ok


def main():
    variants = [1, 2]
    from gpu_raytracer_b200 import _lib as lib
    for n in (1, 2, 3, 5, 31, 127, 128, 129, 255, 256, 257, 383, 1000, 4097, 65536, 100001):
        tris = g.gen_scene(lib.RT_SCENE_SOUP, n, 0, 7)
        for kb in (30, 63):
            print(json.dumps(dict(check="soup", n=n, key_bits=kb, same=check(tris, kb, variants))), flush=True)
    # duplicates: all triangles identical (equal keys everywhere, ties broken by index)
    one = g.gen_scene(lib.RT_SCENE_SOUP, 1, 0, 7)
    dup = one.repeat(3000, 1).contiguous()
    print(json.dumps(dict(check="dup3000", same=check(dup, 30, variants))), flush=True)
    for name in sys.argv[1:] or ["cornell", "soup1m", "flake10m"]:
        wl = WORKLOADS[name]
        tris = g.gen_scene(*wl.scene)
        print(json.dumps(dict(check=name, same=check(tris, wl.key_bits, variants))), flush=True)
        for rnd in range(2):
            for v in [0] + variants:
                g.set_param("build.tree_variant", v)
                time_build(tris, wl.key_bits, 3)
                best, med = time_build(tris, wl.key_bits, 20)
                with g.Bvh.build(tris, key_bits=wl.key_bits, timings=True) as b:
                    st = b.build_times()
                print(json.dumps(dict(scene=name, variant=v, round=rnd, best_ms=round(best, 4),
                                      median_ms=round(med, 4),
                                      tree_ms=round(st.get("topology", 0) + st.get("refit", 0), 4))), flush=True)
        g.set_param("build.tree_variant", 0)
        del tris
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
