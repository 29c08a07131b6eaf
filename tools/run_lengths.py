import sys, numpy as np, torch
sys.path.insert(0, "/root/repo")
import gpu_raytracer_b200 as g
from tools.probe import scene
for name in sys.argv[1:]:
    tris = scene(name)
    with g.Bvh.build(tris, key_bits=63, keep_keys=True) as b:
        keys, idx = b.read_sorted()
    k = torch.from_numpy(keys.view(np.int64)).cuda() >> 33
    del keys, idx
    _, counts = torch.unique_consecutive(k, return_counts=True)
    c = counts.float()
    q = torch.tensor([0.5, 0.9, 0.99, 0.999, 0.9999], device="cuda")
    srt = torch.sort(counts).values
    qs = [int(srt[int(x * (len(srt) - 1))]) for x in q.tolist()]
    print(name, "n", k.numel(), "runs", counts.numel(), "max", int(counts.max()), "quantiles", qs,
          "sum L^2", float((c * c).sum()), "in runs>1", int(counts[counts > 1].sum()),
          ">32:", int((counts > 32).sum()), ">64:", int((counts > 64).sum()), ">256:", int((counts > 256).sum()), flush=True)
    del k, counts, srt, tris
    torch.cuda.empty_cache()
