import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpu_raytracer_b200 as g
from tools.probe import scene
tris = scene(sys.argv[1] if len(sys.argv) > 1 else "terrain100m")
g.set_param("sort.hybrid", 1)
for _ in range(2):
    with g.Bvh.build(tris, key_bits=63) as b:
        pass
torch.cuda.synchronize()
