"""T6 on real GPUs (run under torchrun, one rank per GPU): the concatenated shard
outputs equal the single-GPU output byte for byte, for both BVH strategies
(redundant build, build on rank 0 + NCCL broadcast)."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpu_raytracer_b200 as g  # noqa: E402
from gpu_raytracer_b200 import sharding  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
    n_rays = 1_000_003
    tris = g.gen_scene(g.RT_SCENE_SPHEREFLAKE, 3, 6, 0)
    for strategy in ("redundant", "broadcast"):
        if strategy == "redundant":
            bvh = g.Bvh.build(tris)
        else:
            bvh = sharding.broadcast_bvh(g.Bvh.build(tris) if rank == 0 else None, 0)
        box = bvh.bounds
        parts = [h for _, _, h in sharding.trace_sharded(
            bvh.trace, lambda f, c: g.gen_random_rays(box, 4, f, c), n_rays, rank, world, batch=300_000)]
        local = torch.cat(parts)
        counts = [sharding.shard_range(n_rays, r, world)[1] for r in range(world)]
        assert local.shape[0] == counts[rank]
        # the same shard in one call with RT_TRACE_REORDER: same bytes
        f0, c0 = sharding.shard_range(n_rays, rank, world)
        again = bvh.trace(g.gen_random_rays(box, 4, f0, c0), g.RT_TRACE_REORDER)
        assert torch.equal(again.view(torch.int32), local.view(torch.int32)), "reordered shard differs"
        pad = torch.zeros((max(counts), 4), dtype=torch.float32, device="cuda")
        pad[: local.shape[0]] = local
        outs = [torch.empty_like(pad) for _ in range(world)]
        dist.all_gather(outs, pad)
        full = torch.cat([outs[r][: counts[r]] for r in range(world)])
        if rank == 0:
            single = g.Bvh.build(tris)
            want = single.trace(g.gen_random_rays(box, 4, 0, n_rays))
            same = torch.equal(full.view(torch.int32), want.view(torch.int32))
            hits = int((g.hit_prims(want) >= 0).sum())
            print(f"T6 {strategy}: world={world} rays={n_rays} hits={hits} byte-equal={same}", flush=True)
            assert same and hits > 1000
            single.close()
        bvh.close()
        dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
