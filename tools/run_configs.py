"""Configs 4 and 5 of BASELINE.json on this rank's GPU.

config 4  100M-triangle terrain: 63-bit (and 30-bit) build, per-stage times,
          size-independent checks (sortedness, permutation, sampled tree/box
          invariants) instead of the full CPU oracle.
config 5  10^9 incoherent random rays against the 10M-triangle sphereflake,
          generated on device per batch from (seed, global index); under
          torchrun every rank takes its contiguous range (weak/strong per flag).
"""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import gpu_raytracer_b200 as g  # noqa: E402
from gpu_raytracer_b200 import sharding  # noqa: E402
from gpu_raytracer_b200.configs import WORKLOADS  # noqa: E402

PEAK = 6557.4
BYTES = {30: 388, 63: 528}


def emit(**kw):
    print(json.dumps(kw), flush=True)


def config4():
    wl = WORKLOADS["terrain100m"]
    tris = g.gen_scene(*wl.scene)
    n = tris.shape[0]
    for key_bits in (63, 30):
        best = None
        for _ in range(3):
            with g.Bvh.build(tris, key_bits=key_bits, timings=True) as b:
                t = b.build_times()
            if best is None or t["total"] < best["total"]:
                best = t
        emit(config=4, scene=wl.name, n_tris=n, key_bits=key_bits, mtri_s=n / best["total"] / 1e3,
             roofline_frac=BYTES[key_bits] * n / best["total"] / 1e6 / PEAK, **best)
    # properties at full size (63-bit)
    with g.Bvh.build(tris, key_bits=63, keep_keys=True) as b:
        keys, idx = b.read_sorted()
        sorted_ok = bool((keys[1:] >= keys[:-1]).all())
        perm_ok = bool(np.array_equal(np.sort(idx), np.arange(n, dtype=np.uint32)))
        dup = int((keys[1:] == keys[:-1]).sum())
        left, right, parent = b.read_topology()
        ar = np.arange(n - 1, dtype=np.uint32)
        topo_ok = bool(np.array_equal(parent[left], ar) and np.array_equal(parent[right], ar)
                       and parent[0] == 0xFFFFFFFF)
        kids = np.concatenate([left, right])
        once_ok = bool(np.array_equal(np.sort(kids), np.arange(1, 2 * n - 1, dtype=np.uint32)))
        boxes = b.read_boxes()
        union_ok = bool(np.array_equal(boxes[: n - 1, :3], np.minimum(boxes[left, :3], boxes[right, :3]))
                        and np.array_equal(boxes[: n - 1, 3:], np.maximum(boxes[left, 3:], boxes[right, 3:])))
        emit(config=4, check="properties", sorted=sorted_ok, permutation=perm_ok, duplicate_keys=dup,
             parents_consistent=topo_ok, every_node_child_once=once_ok, boxes_are_unions=union_ok,
             root_box=[float(x) for x in boxes[0]])
        assert sorted_ok and perm_ok and topo_ok and once_ok and union_ok


def config5(total, reorder=False):
    flags = g.RT_TRACE_REORDER if reorder else 0
    for env, name in (("RT_LEAF_MIN", "trace.leaf_min"), ("RT_REFILL_MIN", "trace.refill_min"), ("RT_CHUNK", "trace.chunk")):
        if os.environ.get(env):
            g.set_param(name, int(os.environ[env]))
    if os.environ.get("RT_SH_STACK"):
        g.set_param("trace.sh_stack", int(os.environ["RT_SH_STACK"]))
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
        dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
    wl = WORKLOADS["random1b"]
    tris = g.gen_scene(*wl.scene)
    bvh = g.Bvh.build(tris)
    box = bvh.bounds
    first, count = sharding.shard_range(total, rank, world)
    batch = 1 << 25
    rays = torch.empty((batch, 8), dtype=torch.float32, device="cuda")
    hits = torch.empty((batch, 4), dtype=torch.float32, device="cuda")
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    trace_ms, nhit, done = 0.0, 0, 0
    # warm-up
    g.gen_random_rays(box, wl.random_seed, first, min(batch, count), out=rays)
    bvh.trace_async(rays[: min(batch, count)], hits, flags)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    while done < count:
        c = min(batch, count - done)
        g.gen_random_rays(box, wl.random_seed, first + done, c, out=rays)
        e0.record()
        bvh.trace_async(rays[:c], hits[:c], flags)
        e1.record()
        torch.cuda.synchronize()
        trace_ms += e0.elapsed_time(e1)
        nhit += int((g.hit_prims(hits[:c]) >= 0).sum())
        done += c
    wall = time.perf_counter() - t0
    vals = torch.tensor([trace_ms, wall], dtype=torch.float64, device="cuda")
    tot = torch.tensor([float(nhit)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(vals, op=dist.ReduceOp.MAX)
        dist.all_reduce(tot)
    if rank == 0:
        emit(config=5, scene=wl.name, reorder=bool(reorder), n_tris=int(tris.shape[0]), rays=total, n_gpus=world,
             trace_ms=float(vals[0]), mrays_s=total / float(vals[0]) / 1e3,
             wall_s_incl_generation=float(vals[1]), hit_fraction=float(tot[0]) / total)
    bvh.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    which = sys.argv[1] if len(sys.argv) > 1 else "4"
    if which == "4":
        config4()
    else:
        config5(int(float(sys.argv[2])) if len(sys.argv) > 2 else 10**9,
                reorder=len(sys.argv) > 3 and sys.argv[3] == "reorder")
