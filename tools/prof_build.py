"""Builds one scene a few times (development tool: the command ncu wraps)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpu_raytracer_b200 as g  # noqa: E402
from tools.probe import scene  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "flake10m"
key_bits = int(sys.argv[2]) if len(sys.argv) > 2 else 30
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
wide = len(sys.argv) > 4 and sys.argv[4] == "wide"
tris = scene(name)
for _ in range(reps):
    with g.Bvh.build(tris, key_bits=key_bits, timings=True, wide=wide) as b:
        print(b.build_times(), flush=True)
torch.cuda.synchronize()
