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


# ---- sharded BUILD (§8f.4): every rank holds a part of the scene, hits are exchanged ------------
def _scene_worker(rank, world, port, n_rays, out_dir):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import torch
        import oracle as O
        from gpu_raytracer_b200 import sharding
        O.set_threads(2)
        tris = O.gen_scene(O.SCENE_SOUP, 4000, 0, seed=1)
        box = O.OracleBvh(tris).bounds
        rays = O.gen_random_rays(box, 4, 0, n_rays)                 # the same batch on every rank
        t0, tc = sharding.tri_range(tris.shape[0], rank, world)
        part = O.OracleBvh(tris[t0:t0 + tc])                         # this rank's part of the scene
        for flags in (0, 1):
            local = torch.from_numpy(part.trace(rays, flags))       # all rays against this part
            counts = sharding.exchange_layout(n_rays, world)
            first, mine = sharding.shard_range(n_rays, rank, world)
            recv = torch.empty((world * mine, 4), dtype=torch.float32)
            # gloo has no all_to_all: the same exchange as ShardedScene.trace, as one gather per owner
            for owner in range(world):
                f, c = sharding.shard_range(n_rays, owner, world)
                pieces = [torch.empty((c, 4), dtype=torch.float32) for _ in range(world)] if rank == owner else None
                dist.gather(local[f:f + c].contiguous(), pieces, dst=owner)
                if rank == owner:
                    recv = torch.cat(pieces)
            bases = [sharding.tri_range(tris.shape[0], r, world)[0] for r in range(world)]
            merged = sharding.merge_hits_host([recv[r * mine:(r + 1) * mine].numpy() for r in range(world)],
                                              bases, any_hit=bool(flags))
            np.save(os.path.join(out_dir, f"m{flags}_{rank}.npy"), merged)
    finally:
        dist.destroy_process_group()


def test_sharded_scene_two_ranks_equal_one_bvh(tmp_path, oracle):
    from gpu_raytracer_b200 import sharding
    n_rays = 10007
    mp.spawn(_scene_worker, args=(2, _free_port(), n_rays, str(tmp_path)), nprocs=2, join=True)
    tris = oracle.gen_scene(oracle.SCENE_SOUP, 4000, 0, seed=1)
    bvh = oracle.OracleBvh(tris)
    rays = oracle.gen_random_rays(bvh.bounds, 4, 0, n_rays)
    for flags in (0, 1):
        got = np.concatenate([np.load(str(tmp_path / f"m{flags}_{r}.npy")) for r in range(2)])
        want = bvh.trace(rays, flags)
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), flags
    assert [sharding.tri_range(10, r, 3) for r in range(3)] == [(0, 3), (3, 3), (6, 4)]


# ---- the scene crosses PCIe once: every rank uploads 1/world of it, one all-gather ----------------
def _upload_worker(rank, world, port, n_tris, out_dir):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import torch
        import oracle as O
        from gpu_raytracer_b200 import sharding
        tris = torch.from_numpy(O.gen_scene(O.SCENE_SOUP, n_tris, 0, seed=1))
        full = sharding.upload_scene(tris, rank, world, device="cpu")
        np.save(os.path.join(out_dir, f"t{n_tris}_{rank}.npy"), full.numpy())
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n_tris", [4000, 4001])     # equal shares, ragged shares
def test_upload_scene_gathers_the_whole_array_on_every_rank(tmp_path, oracle, n_tris):
    mp.spawn(_upload_worker, args=(2, _free_port(), n_tris, str(tmp_path)), nprocs=2, join=True)
    want = oracle.gen_scene(oracle.SCENE_SOUP, n_tris, 0, seed=1)
    for r in range(2):
        got = np.load(str(tmp_path / f"t{n_tris}_{r}.npy"))
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32)), r


def test_merge_of_parts_equals_the_whole(oracle):
    """The property the sharded build rests on: a ray's hit is min over ALL triangles by (t, prim),
    so the parts of any partition of the triangle array merge to the single-BVH hits exactly."""
    from gpu_raytracer_b200 import sharding
    for kind, p0, p1, seed in [(0, 0, 0, 0), (1, 3000, 0, 5), (2, 2, 3, 0)]:
        tris = oracle.gen_scene(kind, p0, p1, seed)
        whole = oracle.OracleBvh(tris)
        rays = oracle.gen_random_rays(whole.bounds, 4, 0, 30000)
        for world in (2, 3, 7):
            cuts = [sharding.tri_range(tris.shape[0], r, world) for r in range(world)]
            for flags in (0, oracle.TRACE_ANY, oracle.TRACE_WATERTIGHT):
                parts = [oracle.OracleBvh(tris[f:f + c]).trace(rays, flags) for f, c in cuts]
                got = sharding.merge_hits_host(parts, [f for f, _ in cuts], any_hit=bool(flags & 1))
                assert np.array_equal(got.view(np.uint32), whole.trace(rays, flags).view(np.uint32)), (kind, world, flags)
