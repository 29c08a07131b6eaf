"""T6 on CPU: the N>1 host logic with world_size 2 over gloo.

The sharding code is the product's (gpu_raytracer_b200.sharding); the tracer
plugged into it here is the oracle, because this container has no GPU.  The
check is the one the GPU run repeats: concatenated shard outputs == the
single-process output, byte for byte, and rays depend on (seed, index) only.
"""
import os
import socket
import sys

import numpy as np
import pytest
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n_rays, out_path):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import oracle as O
        from gpu_raytracer_b200 import sharding
        O.set_threads(2)
        tris = O.gen_scene(O.SCENE_SOUP, 4000, 0, seed=1)
        bvh = O.OracleBvh(tris)  # "redundant build" strategy: every rank builds
        parts = []
        for first, count, hits in sharding.trace_sharded(
                bvh.trace, lambda f, c: O.gen_random_rays(bvh.bounds, 4, f, c), n_rays, rank, world,
                batch=1500):
            parts.append(hits)
        local = np.concatenate(parts) if parts else np.zeros((0, 4), dtype=np.float32)
        assert local.shape[0] == sharding.shard_range(n_rays, rank, world)[1]
        full = sharding.gather_hits(local, n_rays, rank, world)
        if rank == 0:
            np.save(out_path, full)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n_rays", [10007])
def test_two_ranks_equal_one(tmp_path, oracle, n_rays):
    out = str(tmp_path / "hits.npy")
    mp.spawn(_worker, args=(2, _free_port(), n_rays, out), nprocs=2, join=True)
    got = np.load(out)
    tris = oracle.gen_scene(oracle.SCENE_SOUP, 4000, 0, seed=1)
    bvh = oracle.OracleBvh(tris)
    want = bvh.trace(oracle.gen_random_rays(bvh.bounds, 4, 0, n_rays))
    assert got.shape == want.shape
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    assert (oracle.hit_prims(want) >= 0).sum() > 100
