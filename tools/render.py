"""Renders a scene to a PPM or PNG file (by the name's ending) through the public API (SURVEY.md
§8f.1): build_bvh, then the whole frame in one call (rt_render_camera_ao: primary rays from the camera,
4 ambient-occlusion rays per pixel, shading), image.write_ppm / write_png.

  python tools/render.py cornell|soup1m|flake10m|<file.obj>|<file.tri> [out.ppm|out.png] [width height]
"""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpu_raytracer_b200 as g  # noqa: E402
from gpu_raytracer_b200.configs import WORKLOADS  # noqa: E402


def main():
    what = sys.argv[1] if len(sys.argv) > 1 else "cornell"
    out = sys.argv[2] if len(sys.argv) > 2 else "gpurun_out/render.ppm"
    if what in WORKLOADS:
        wl = WORKLOADS[what]
        tris = g.gen_scene(*wl.scene)
        w, h = wl.width, wl.height
        cam = wl.camera()
    else:
        host = g.load_obj(what) if what.endswith(".obj") else g.load_tris(what)
        tris = torch.from_numpy(host).cuda()
        w, h = 1280, 720
        cam = None
    if len(sys.argv) > 4:
        w, h = int(sys.argv[3]), int(sys.argv[4])
        cam = WORKLOADS[what].camera(w, h) if what in WORKLOADS else None
    t0 = time.perf_counter()
    with g.Bvh.build(tris) as bvh:
        box = bvh.bounds
        if cam is None:
            cam = g.camera.frame_box(box, w, h)
        radius = float(np.float32(0.05) * np.float32(np.linalg.norm(box[3:] - box[:3])))
        hits, rgb = bvh.render_camera_ao(tris, cam, w, h, 4, 2, radius, device="cuda")
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    os.makedirs(os.path.dirname(os.path.abspath(out)), exist_ok=True)
    (g.image.write_png if out.lower().endswith(".png") else g.image.write_ppm)(out, rgb, w, h)
    cover = float((g.hit_prims(hits) >= 0).float().mean())
    print(f"{what}: {tris.shape[0]} triangles, {w}x{h}, coverage {cover:.3f}, "
          f"build+trace+shade {dt * 1e3:.1f} ms -> {out}")


if __name__ == "__main__":
    main()
