import os, sys, torch
sys.path.insert(0, "/root/repo")
import gpu_raytracer_b200 as g
from tools.probe import scene
tris = scene("flake10m")
g.set_param("build.tree_variant", int(sys.argv[1]))
for _ in range(2):
    with g.Bvh.build(tris, key_bits=30) as b:
        pass
torch.cuda.synchronize()
