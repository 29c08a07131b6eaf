"""Generates tests/golden/*.npz from the CPU oracle.

The reference holds no golden vectors, so these fixtures pin the
oracle's *own* outputs at the commit that produced them: a later change to the
oracle or to the kernels that alters any byte fails tests/test_golden.py.  They
travel to the GPU box, where the CUDA path is compared with them directly.

    python tests/golden/make_golden.py        # rewrites the fixtures
    python tests/golden/make_golden.py frame  # only cornell_96_frame.json (§8f.1 image)
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import oracle as O  # noqa: E402
from gpu_raytracer_b200 import camera  # noqa: E402


def digest(a):
    return np.frombuffer(hashlib.sha256(np.ascontiguousarray(a).tobytes()).digest(), dtype=np.uint8)


def build_fixture(name, tris, key_bits, rays, extra=None):
    b = O.OracleBvh(tris, key_bits)
    hits = b.trace(rays)
    occ = b.trace(rays, O.TRACE_ANY)
    wt_hits = b.trace(rays, O.TRACE_WATERTIGHT)             # A7 option
    wide_boxes, wide_desc = b.collapse4(2)                  # A5w, default leaf size
    out = dict(wt_hits=wt_hits, wide_desc=wide_desc, wide_boxes_sha=digest(wide_boxes),
               key_bits=np.int32(key_bits), n=np.int64(b.n), bounds=b.bounds,
               tris_sha=digest(tris), keys=b.keys, idx=b.idx, left=b.left, right=b.right,
               parent=b.parent, boxes_sha=digest(b.boxes), boxes_head=b.boxes[:64],
               rays_sha=digest(rays), hits=hits, any_prim=O.hit_prims(occ))
    out.update(extra or {})
    np.savez_compressed(os.path.join(HERE, name + ".npz"), **out)
    print(name, b.n, "tris", rays.shape[0], "rays", int((O.hit_prims(hits) >= 0).sum()), "hits")


def cornell_frame(width=96, height=96, spp=4, seed=2):
    """§8(f).1: Cornell box, primary rays -> AO rays -> any-hit -> shade, all by the oracle.
    Returns (rgb float32 [w*h, 3], radius)."""
    cam = camera.cornell_camera(width, height)
    tris = O.gen_scene(O.SCENE_CORNELL)
    b = O.OracleBvh(tris)
    rays = O.gen_primary_rays(cam, width, height)
    hits = b.trace(rays)
    radius = np.float32(0.05) * np.float32(np.linalg.norm(b.bounds[3:] - b.bounds[:3]))
    ao = O.gen_ao_rays(tris, rays, hits, spp, seed, float(radius))
    occ = b.trace(ao, O.TRACE_ANY)
    return O.shade(tris, rays, hits, occ, spp), float(radius)


def frame_fixture():
    import json
    import tempfile
    from gpu_raytracer_b200 import image
    rgb, radius = cornell_frame()
    with tempfile.TemporaryDirectory() as d:
        path = os.path.join(d, "f.ppm")
        image.write_ppm(path, rgb, 96, 96)
        ppm = open(path, "rb").read()
    out = {"width": 96, "height": 96, "spp": 4, "seed": 2, "radius": radius,
           "rgb_sha256": hashlib.sha256(rgb.tobytes()).hexdigest(),
           "ppm_bytes": len(ppm), "ppm_sha256": hashlib.sha256(ppm).hexdigest(),
           "mean": float(rgb.mean())}
    with open(os.path.join(HERE, "cornell_96_frame.json"), "w") as f:
        json.dump(out, f, indent=1)
    print("cornell_96_frame", out)


def main():
    if sys.argv[1:] == ["frame"]:
        frame_fixture()
        return
    frame_fixture()
    # config 1, reduced image: Cornell box, 96 x 96 primary rays
    cam = camera.cornell_camera(96, 96)
    tris = O.gen_scene(O.SCENE_CORNELL)
    build_fixture("cornell_96", tris, 30, O.gen_primary_rays(cam, 96, 96), dict(cam=cam))
    # soup, both key widths, random rays
    tris = O.gen_scene(O.SCENE_SOUP, 3000, 0, seed=1)
    v = tris.reshape(-1, 4)[:, :3]
    box = np.concatenate([v.min(0), v.max(0)])
    rays = O.gen_random_rays(box, 4, 0, 4096)
    build_fixture("soup3000_k30", tris, 30, rays, dict(box=box))
    build_fixture("soup3000_k63", tris, 63, rays, dict(box=box))
    # sphereflake and terrain slices
    tris = O.gen_scene(O.SCENE_SPHEREFLAKE, 2, 3)
    v = tris.reshape(-1, 4)[:, :3]
    box = np.concatenate([v.min(0), v.max(0)])
    build_fixture("flake_d2_m3", tris, 30, O.gen_random_rays(box, 4, 0, 4096), dict(box=box))
    tris = O.gen_scene(O.SCENE_TERRAIN, 48, 16, seed=3)
    v = tris.reshape(-1, 4)[:, :3]
    box = np.concatenate([v.min(0), v.max(0)])
    build_fixture("terrain_48", tris, 63, O.gen_random_rays(box, 4, 0, 4096), dict(box=box))


if __name__ == "__main__":
    main()
