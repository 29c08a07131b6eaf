"""Multi-GPU sharding of ray batches.

One process per GPU (torch.distributed).  Rays are independent, so the batch is
cut into contiguous ranges, one per rank, with no collective on the data path;
each rank generates (or receives) only its range, and ray i depends on (seed, i)
alone, so results do not depend on the number of GPUs.  The BVH is either built
redundantly on every rank or built on rank 0 and broadcast (one NCCL broadcast
of nodes + triangles + header over NVLink); ``bench.py`` measures both.
"""
from __future__ import annotations

from typing import Callable, Optional, Tuple

import numpy as np


def shard_range(n: int, rank: int, world: int, align: int = 32) -> Tuple[int, int]:
    """(first, count) of rank's contiguous range of [0, n).

    Ranges are cut at multiples of ``align`` rays (a warp's fetch granule) except
    for the last one; they tile [0, n) exactly and differ by less than two granules."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    units = (n + align - 1) // align
    base, rem = divmod(units, world)
    u0 = rank * base + min(rank, rem)
    u1 = u0 + base + (1 if rank < rem else 0)
    first = min(u0 * align, n)
    last = min(u1 * align, n)
    return first, last - first


class _DevBuf:
    """Exposes a raw device pointer to torch through __cuda_array_interface__."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {
            "shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 3, "strides": None}


def _as_tensor(ptr: int, nbytes: int):
    import torch
    return torch.as_tensor(_DevBuf(ptr, nbytes), device="cuda")


def broadcast_bvh(bvh, src: int = 0, group=None):
    """Broadcasts a built BVH from ``src`` to every rank; returns the local handle.

    ``bvh`` is the built handle on ``src`` and ignored (may be None) elsewhere."""
    import torch
    import torch.distributed as dist
    from .api import Bvh

    rank = dist.get_rank(group)
    meta = torch.zeros(2, dtype=torch.int64, device="cuda")
    if rank == src:
        info = bvh.info()
        meta[0], meta[1] = int(info.n_tris), int(info.key_bits)
    dist.broadcast(meta, src=src, group=group)
    n_tris, key_bits = int(meta[0].item()), int(meta[1].item())
    if rank != src:
        bvh = Bvh.alloc_empty(n_tris, key_bits)
    for ptr, nbytes in bvh.device_buffers():
        if nbytes:
            dist.broadcast(_as_tensor(ptr, nbytes), src=src, group=group)
    torch.cuda.synchronize()
    if rank != src:
        bvh.commit()
    return bvh


def upload_scene(host_tris, rank: int, world: int, group=None, device=None):
    """The whole triangle array [n, 12] on every rank's device, with the array crossing PCIe ONCE
    in total: every rank uploads rows ``tri_range(n, rank, world)`` of the (identical) host array
    and one all-gather (NCCL over NVLink on GPUs; gloo on CPU tensors in the tests) hands every
    rank the rest.  Against every rank uploading everything, the ranks of one host stop competing
    for its PCIe bandwidth (8 x 479 MB at 10M triangles).  Returns a tensor on ``device``."""
    import torch
    import torch.distributed as dist

    n = int(host_tris.shape[0])
    device = torch.device(device if device is not None else ("cuda" if torch.cuda.is_available() else "cpu"))
    if world == 1:
        return host_tris.to(device, non_blocking=True)
    counts = [n * (r + 1) // world - n * r // world for r in range(world)]
    first = n * rank // world
    full = torch.empty((n, 12), dtype=torch.float32, device=device)
    mine = host_tris[first:first + counts[rank]].to(device, non_blocking=True)
    if len(set(counts)) == 1:
        dist.all_gather_into_tensor(full, mine.contiguous(), group=group)
    else:  # ragged: equal-size padded pieces, trimmed into place
        mx = max(counts)
        pad = torch.zeros((mx, 12), dtype=torch.float32, device=device)
        pad[:counts[rank]] = mine
        pieces = [torch.empty_like(pad) for _ in range(world)]
        dist.all_gather(pieces, pad, group=group)
        at = 0
        for r in range(world):
            full[at:at + counts[r]] = pieces[r][:counts[r]]
            at += counts[r]
    return full


def trace_sharded(trace_fn: Callable, make_rays: Callable, n_rays: int, rank: int, world: int,
                  batch: Optional[int] = None):
    """Traces this rank's range of a global batch of ``n_rays`` rays.

    make_rays(first, count) returns the rays of the global range; trace_fn(rays)
    returns their hits.  Yields (first, count, hits) per batch so the caller
    decides whether hits stay on the device, go to the host or are reduced."""
    first, count = shard_range(n_rays, rank, world)
    batch = batch or count or 1
    done = 0
    while done < count:
        c = min(batch, count - done)
        rays = make_rays(first + done, c)
        yield first + done, c, trace_fn(rays)
        done += c


def gather_hits(local_hits: np.ndarray, n_rays: int, rank: int, world: int, group=None) -> np.ndarray:
    """All ranks' hit ranges concatenated in ray order (host arrays; used by the
    T6 test: concatenated shard outputs == single-GPU output, byte for byte)."""
    import torch
    import torch.distributed as dist

    counts = [shard_range(n_rays, r, world)[1] for r in range(world)]
    mx = max(counts) if counts else 0
    pad = np.zeros((mx, 4), dtype=np.float32)
    pad[:local_hits.shape[0]] = local_hits
    mine = torch.from_numpy(pad)
    outs = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(outs, mine, group=group)
    return np.concatenate([outs[r].numpy()[:counts[r]] for r in range(world)], axis=0)


# ---------------------------------------------------------------------------
# Sharded BUILD: the scene is cut, the hits are exchanged
# ---------------------------------------------------------------------------
def tri_range(n_tris: int, rank: int, world: int) -> Tuple[int, int]:
    """(first, count) of rank's contiguous range of the triangle array (build_bvh_sharded's cut)."""
    first = n_tris * rank // world
    return first, n_tris * (rank + 1) // world - first


def merge_hits_host(parts, prim_base, any_hit: bool = False) -> np.ndarray:
    """numpy statement of rt_merge_hits (tests; the product path is the CUDA kernel): per ray
    the minimum over the parts by (t, prim + base); any-hit: occluded if any part is."""
    parts = [np.ascontiguousarray(p, dtype=np.float32).reshape(-1, 4) for p in parts]
    n = parts[0].shape[0]
    out = np.zeros((n, 4), dtype=np.float32)
    out[:, 0] = np.inf
    best_prim = np.full(n, -1, dtype=np.int64)
    for p, base in zip(parts, prim_base):
        prim = np.ascontiguousarray(p[:, 3]).view(np.int32).astype(np.int64)
        hit = prim >= 0
        if any_hit:
            best_prim[hit] = 0x7FFFFFFF
            continue
        gp = prim + int(base)
        better = hit & ((best_prim < 0) | (p[:, 0] < out[:, 0]) | ((p[:, 0] == out[:, 0]) & (gp < best_prim)))
        out[better, :3] = p[better, :3]
        best_prim[better] = gp[better]
    if any_hit:
        out[best_prim >= 0, 0] = 0.0
    out[:, 3] = best_prim.astype(np.int32).view(np.float32)
    return out


def exchange_layout(n_rays: int, world: int):
    """Send/receive counts (in rays) of the all-to-all that turns "every rank holds the hits of
    ALL rays against ITS part" into "every rank holds ALL parts' hits of ITS ray range"."""
    counts = [shard_range(n_rays, r, world)[1] for r in range(world)]
    return counts


class ShardedScene:
    """One part of a scene per rank (one process per GPU).  Every rank traces the whole ray
    batch against its part; one all-to-all moves, for every ray range, the parts' hit records
    to the rank that owns the range; rt_merge_hits reduces them by (t, prim).  The result is
    sharded by rays like everything else here, and equals the single-BVH hits bit for bit."""

    def __init__(self, bvh, first_tri: int, n_tris_total: int, group=None):
        self.bvh, self.first_tri, self.n_tris_total, self.group = bvh, first_tri, n_tris_total, group

    @classmethod
    def build(cls, tris_local, first_tri: int, n_tris_total: int, key_bits: int = 30, group=None):
        from .api import Bvh
        return cls(Bvh.build(tris_local, key_bits=key_bits), first_tri, n_tris_total, group)

    def trace(self, rays, flags: int = 0):
        """rays: CUDA tensor [n, 8], the SAME batch on every rank.  Returns (first, hits) of this
        rank's ray range, merged over all parts."""
        import torch
        import torch.distributed as dist
        from . import api
        rank, world = dist.get_rank(self.group), dist.get_world_size(self.group)
        n = int(rays.shape[0])
        local = self.bvh.trace(rays, flags)                       # hits of all rays against this part
        counts = exchange_layout(n, world)
        first, mine = shard_range(n, rank, world)
        recv = torch.empty((world * mine, 4), dtype=torch.float32, device=rays.device)
        dist.all_to_all_single(recv, local, output_split_sizes=[mine] * world,
                               input_split_sizes=counts, group=self.group)
        bases = [tri_range(self.n_tris_total, r, world)[0] for r in range(world)]
        parts = [recv[r * mine:(r + 1) * mine] for r in range(world)]
        return first, api.merge_hits(parts, bases, flags & 1) if mine else recv[:0]

    def close(self):
        self.bvh.close()
