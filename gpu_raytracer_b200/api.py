"""Host-side mirror of the scene-in / hits-out interface (BASELINE.json:5).

Thin Python over the C ABI in include/gpu_raytracer.h: ``Bvh.build`` →
``build_bvh``, ``Bvh.trace`` → ``trace_rays``, ``Bvh.close`` → ``free_bvh``.
Buffers may be numpy arrays (host) or torch tensors (host, pinned or CUDA); the
library decides per pointer whether to stage or to use memory in place.  torch
is only a device-memory allocator here.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import numpy as np

from . import _lib as L

try:  # torch is plumbing (device memory, streams), not a dependency of the ABI
    import torch
except Exception:  # pragma: no cover
    torch = None


def _is_tensor(x) -> bool:
    return torch is not None and isinstance(x, torch.Tensor)


def _ptr(x, dtype=np.float32) -> int:
    """Raw address of a contiguous numpy array or torch tensor."""
    if _is_tensor(x):
        if not x.is_contiguous():
            raise ValueError("tensor must be contiguous")
        return x.data_ptr()
    if isinstance(x, np.ndarray):
        if not x.flags["C_CONTIGUOUS"]:
            raise ValueError("array must be C-contiguous")
        if dtype is not None and x.dtype != dtype:
            raise ValueError(f"array must be {dtype}, got {x.dtype}")
        return x.ctypes.data
    raise TypeError(f"expected numpy array or torch tensor, got {type(x)}")


def _rows(x, width: int) -> int:
    n = int(np.prod(x.shape))
    if n % width:
        raise ValueError(f"buffer size {n} is not a multiple of {width}")
    return n // width


def _empty_like_kind(ref, shape, device: Optional[str]):
    """Output buffer of the same kind as ``ref`` (or on ``device`` when given)."""
    if device is not None or _is_tensor(ref):
        dev = device if device is not None else ref.device
        return torch.empty(shape, dtype=torch.float32, device=dev)
    return np.empty(shape, dtype=np.float32)


# ---------------------------------------------------------------------------
# generators
# ---------------------------------------------------------------------------
def scene_size(kind: int, p0: int = 0, p1: int = 0) -> int:
    return int(L.lib().rt_scene_size(kind, p0, p1))


def gen_scene(kind: int, p0: int = 0, p1: int = 0, seed: int = 0, first: int = 0,
              count: Optional[int] = None, device: Optional[str] = "cuda"):
    """Triangles [first, first+count) of a synthetic scene as float32 [count, 12].

    device="cuda" returns a CUDA tensor (generated in place, never on the host);
    device=None returns a numpy array (generated on the GPU, copied back)."""
    if count is None:
        count = scene_size(kind, p0, p1) - first
    out = (torch.empty((count, 12), dtype=torch.float32, device=device) if device is not None
           else np.empty((count, 12), dtype=np.float32))
    L.check(L.lib().rt_gen_scene(kind, p0, p1, seed, first, count, _ptr(out)))
    return out


def gen_primary_rays(cam, width: int, height: int, first: int = 0, count: Optional[int] = None,
                     device: Optional[str] = "cuda"):
    cam = np.ascontiguousarray(cam, dtype=np.float32).reshape(12)
    if count is None:
        count = width * height - first
    out = (torch.empty((count, 8), dtype=torch.float32, device=device) if device is not None
           else np.empty((count, 8), dtype=np.float32))
    L.check(L.lib().rt_gen_primary_rays(cam.ctypes.data, width, height, first, count, _ptr(out)))
    return out


def gen_random_rays(box, seed: int, first: int, count: int, device: Optional[str] = "cuda",
                    out=None):
    box = np.ascontiguousarray(box, dtype=np.float32).reshape(6)
    if out is None:
        out = (torch.empty((count, 8), dtype=torch.float32, device=device) if device is not None
               else np.empty((count, 8), dtype=np.float32))
    L.check(L.lib().rt_gen_random_rays(box.ctypes.data, seed, first, count, _ptr(out)))
    return out


def gen_ao_rays(tris, rays, hits, spp: int, seed: int, radius: float, out=None):
    """spp occlusion rays per primary ray; tris (original order) must be on the device."""
    n = _rows(rays, 8)
    if out is None:
        out = _empty_like_kind(rays, (n * spp, 8), None)
    L.check(L.lib().rt_gen_ao_rays(_ptr(tris), _ptr(rays), _ptr(hits), n, spp, seed,
                                   C.c_float(radius), _ptr(out)))
    return out


def shade(tris, rays, hits, ao_hits=None, spp: int = 0):
    n = _rows(rays, 8)
    out = _empty_like_kind(rays, (n, 3), None)
    L.check(L.lib().rt_shade(_ptr(tris), _ptr(rays), _ptr(hits),
                             _ptr(ao_hits) if ao_hits is not None else None, n, spp, _ptr(out)))
    return out


# one slot of bvh_read_wide (include/gpu_raytracer.h)
WIDE_NODE_DTYPE = np.dtype([("lo", np.float32, (3, 4)), ("hi", np.float32, (3, 4)),
                            ("ref", np.uint32, 4), ("n_children", np.uint32), ("pad", np.uint32, 3)])
WIDE_LEAF_BIT, WIDE_EMPTY, WIDE_FIRST_MASK = 0x80000000, 0x7FFFFFFF, (1 << 29) - 1
# one node of bvh_read_wide8 (80 bytes)
WIDE8_NODE_DTYPE = np.dtype([("p", np.float32, 3), ("e", np.uint8, 3), ("imask", np.uint8),
                             ("child_base", np.uint32), ("tri_base", np.uint32), ("meta", np.uint8, 8),
                             ("qlo", np.uint8, (3, 8)), ("qhi", np.uint8, (3, 8))])
# one node of bvh_read_compact (32 bytes): box[c, a] = lo | hi << 16 of child c on axis a
COMPACT_NODE_DTYPE = np.dtype([("box", np.uint32, (2, 3)), ("child", np.int32, 2)])


# ---------------------------------------------------------------------------
# scene files (host code of the library; no GPU needed)
# ---------------------------------------------------------------------------
def _take_host_tris(ptr: C.c_void_p, n: int) -> np.ndarray:
    try:
        if n == 0:
            return np.zeros((0, 12), dtype=np.float32)
        src = np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_float)), shape=(n, 12))
        return src.copy()
    finally:
        L.lib().rt_free_host(ptr)


def load_obj(path: str) -> np.ndarray:
    """Triangles of a Wavefront OBJ file as float32 [n, 12] (rt_load_obj)."""
    ptr, n = C.c_void_p(), C.c_uint64()
    L.check(L.lib().rt_load_obj(str(path).encode(), C.byref(ptr), C.byref(n)))
    return _take_host_tris(ptr, int(n.value))


def save_tris(path: str, tris: np.ndarray) -> None:
    tris = np.ascontiguousarray(tris, dtype=np.float32).reshape(-1, 12)
    L.check(L.lib().rt_save_tris(str(path).encode(), tris.ctypes.data, tris.shape[0]))


def load_tris(path: str) -> np.ndarray:
    ptr, n = C.c_void_p(), C.c_uint64()
    L.check(L.lib().rt_load_tris(str(path).encode(), C.byref(ptr), C.byref(n)))
    return _take_host_tris(ptr, int(n.value))


# ---------------------------------------------------------------------------
# the BVH handle
# ---------------------------------------------------------------------------
class Bvh:
    """An LBVH on one GPU.  Immutable after build; ``trace`` is thread-safe."""

    def __init__(self, handle: int):
        self._h = C.c_void_p(handle)

    # -- build_bvh -----------------------------------------------------------
    @classmethod
    def build(cls, tris, key_bits: int = 30, keep_keys: bool = False, timings: bool = False,
              device: int = -1, stream: int = 0, wide: bool = False, wide8: bool = False,
              compact: bool = False) -> "Bvh":
        n = _rows(tris, 12)
        opts = L.BuildOpts(key_bits=key_bits,
                           flags=(L.RT_BUILD_KEEP_KEYS if keep_keys else 0) |
                                 (L.RT_BUILD_TIMINGS if timings else 0) |
                                 (L.RT_BUILD_WIDE4 if wide else 0) | (L.RT_BUILD_WIDE8 if wide8 else 0) |
                                 (L.RT_BUILD_COMPACT if compact else 0),
                           device=device, reserved=0, stream=stream or None)
        h = C.c_void_p()
        L.check(L.lib().build_bvh(_ptr(tris) if n else None, n, C.byref(opts), C.byref(h)))
        return cls(h.value)

    @classmethod
    def load(cls, path: str) -> "Bvh":
        """A BVH written by ``save`` (bvh_load)."""
        h = C.c_void_p()
        L.check(L.lib().bvh_load(str(path).encode(), C.byref(h)))
        return cls(h.value)

    def save(self, path: str) -> None:
        L.check(L.lib().bvh_save(self._h, str(path).encode()))

    @classmethod
    def alloc_empty(cls, n_tris: int, key_bits: int = 30) -> "Bvh":
        h = C.c_void_p()
        L.check(L.lib().bvh_alloc_empty(n_tris, key_bits, C.byref(h)))
        return cls(h.value)

    # -- trace_rays ------------------------------------------------------------
    def trace(self, rays, flags: int = L.RT_TRACE_CLOSEST, out=None):
        """hits float32 [n, 4]: t, u, v, prim (int32 bits).  Same kind as ``rays``."""
        n = _rows(rays, 8)
        if out is None:
            out = _empty_like_kind(rays, (n, 4), None)
        L.check(L.lib().trace_rays(self._h, _ptr(rays), n, flags, _ptr(out)))
        return out

    def trace_camera(self, cam, width: int, height: int, flags: int = L.RT_TRACE_CLOSEST, out=None,
                     first: int = 0, count: Optional[int] = None, device: Optional[str] = None):
        """Hits of the primary rays of a pinhole image, generated on the device (trace_camera):
        only the hits cross the bus.  ``out``: numpy / pinned tensor / CUDA tensor [count, 4]."""
        cam = np.ascontiguousarray(cam, dtype=np.float32).reshape(12)
        if count is None:
            count = width * height - first
        if out is None:
            out = (torch.empty((count, 4), dtype=torch.float32, device=device) if device is not None
                   else np.empty((count, 4), dtype=np.float32))
        L.check(L.lib().trace_camera(self._h, cam.ctypes.data, width, height, first, count, flags,
                                     _ptr(out)))
        return out

    def render_ao(self, tris, rays, spp: int, seed: int, radius: float, flags: int = 0, hits=None, rgb=None):
        """One frame in one call (rt_render_ao): primary trace, spp AO rays per pixel, their any-hit
        trace and the shaded frame, pipelined in bands.  tris: CUDA tensor, original order.
        Returns (hits [n, 4], rgb [n, 3]) of the same kind as ``rays``."""
        n = _rows(rays, 8)
        pin = _is_tensor(rays) and not rays.is_cuda and rays.is_pinned()
        if hits is None:
            hits = _empty_like_kind(rays, (n, 4), None)
            hits = hits.pin_memory() if pin else hits
        if rgb is None:
            rgb = _empty_like_kind(rays, (n, 3), None)
            rgb = rgb.pin_memory() if pin else rgb
        L.check(L.lib().rt_render_ao(self._h, _ptr(tris), _ptr(rays), n, spp, seed, C.c_float(radius), flags,
                                     _ptr(hits), _ptr(rgb)))
        return hits, rgb

    def render_camera_ao(self, tris, cam, width: int, height: int, spp: int, seed: int, radius: float,
                         flags: int = 0, hits=None, rgb=None, device: Optional[str] = None):
        """The frame of ``cam`` in one call (rt_render_camera_ao): primary rays generated on the
        device, spp AO rays per pixel, the shaded frame.  Returns (hits [n, 4], rgb [n, 3]): numpy
        unless ``device`` or the buffers say otherwise."""
        cam = np.ascontiguousarray(cam, dtype=np.float32).reshape(12)
        n = width * height
        if hits is None:
            hits = _empty_like_kind(cam, (n, 4), device)
        if rgb is None:
            rgb = _empty_like_kind(cam, (n, 3), device)
        L.check(L.lib().rt_render_camera_ao(self._h, _ptr(tris), cam.ctypes.data, width, height, spp, seed,
                                            C.c_float(radius), flags, _ptr(hits), _ptr(rgb)))
        return hits, rgb

    def trace_async(self, rays, out, flags: int = L.RT_TRACE_CLOSEST, stream: int = 0):
        n = _rows(rays, 8)
        L.check(L.lib().trace_rays_async(self._h, _ptr(rays), n, flags, _ptr(out), stream or None))
        return out

    def trace_stats(self, rays, flags: int = L.RT_TRACE_CLOSEST):
        """(hits, nodes_fetched, triangles_tested) for device rays; measurement only."""
        n = _rows(rays, 8)
        out = _empty_like_kind(rays, (n, 4), None)
        cnt = (C.c_uint64 * 2)()
        L.check(L.lib().trace_rays_stats(self._h, _ptr(rays), n, flags, _ptr(out), cnt))
        return out, int(cnt[0]), int(cnt[1])

    def refit(self, tris) -> None:
        L.check(L.lib().refit_bvh(self._h, _ptr(tris)))

    def widen(self) -> None:
        """Collapse to the 4-wide layout (A5w); ``trace`` then walks it."""
        L.check(L.lib().bvh_widen(self._h))

    def widen8(self) -> None:
        """Collapse to the compressed 8-wide layout (needs keep_keys=True); ``trace`` then walks it."""
        L.check(L.lib().bvh_widen8(self._h))

    def compact(self) -> None:
        """Pack the binary tree into 32-byte nodes (bvh_compact); ``trace`` then walks those."""
        L.check(L.lib().bvh_compact(self._h))

    def read_compact(self):
        """(nodes, grid6): the 32-byte nodes as a structured array and the grid (origin xyz, step xyz)."""
        nodes = np.zeros(max(self.n_tris - 1, 0), dtype=COMPACT_NODE_DTYPE)
        grid = np.zeros(6, dtype=np.float32)
        L.check(L.lib().bvh_read_compact(self._h, nodes.ctypes.data if len(nodes) else None, grid.ctypes.data))
        return nodes, grid

    def read_wide8(self):
        """(nodes, tris): the 8-wide layout as a structured array and the triangles in its order."""
        cnt = C.c_uint32()
        L.check(L.lib().bvh_read_wide8(self._h, None, None, C.byref(cnt)))
        nodes = np.zeros(int(cnt.value), dtype=WIDE8_NODE_DTYPE)
        tris = np.zeros((self.n_tris, 12), dtype=np.float32)
        L.check(L.lib().bvh_read_wide8(self._h, nodes.ctypes.data, tris.ctypes.data, C.byref(cnt)))
        return nodes, tris

    def read_wide(self):
        """(nodes, root_ref): the BVH4 as a structured array with n-1 slots."""
        n = self.n_tris
        nodes = np.zeros(max(n - 1, 0), dtype=WIDE_NODE_DTYPE)
        root = C.c_uint32()
        L.check(L.lib().bvh_read_wide(self._h, nodes.ctypes.data if n > 1 else None,
                                      C.byref(root)))
        return nodes, int(root.value)

    # -- accessors ---------------------------------------------------------------
    def info(self) -> L.BvhInfo:
        info = L.BvhInfo()
        L.check(L.lib().bvh_info(self._h, C.byref(info)))
        return info

    @property
    def n_tris(self) -> int:
        return int(self.info().n_tris)

    @property
    def bounds(self) -> np.ndarray:
        return np.array(list(self.info().bounds), dtype=np.float32)

    def read_sorted(self, keys: bool = True):
        n = self.n_tris
        k = np.empty(n, dtype=np.uint64) if keys else None
        idx = np.empty(n, dtype=np.uint32)
        L.check(L.lib().bvh_read_sorted(self._h, k.ctypes.data if keys else None, idx.ctypes.data))
        return k, idx

    def read_topology(self):
        n = self.n_tris
        left = np.empty(max(n - 1, 0), dtype=np.uint32)
        right = np.empty(max(n - 1, 0), dtype=np.uint32)
        parent = np.empty(max(2 * n - 1, 0), dtype=np.uint32)
        L.check(L.lib().bvh_read_topology(self._h, left.ctypes.data, right.ctypes.data,
                                          parent.ctypes.data))
        return left, right, parent

    def read_boxes(self) -> np.ndarray:
        n = self.n_tris
        boxes = np.empty((max(2 * n - 1, 0), 6), dtype=np.float32)
        L.check(L.lib().bvh_read_boxes(self._h, boxes.ctypes.data))
        return boxes

    def build_times(self) -> dict:
        ms = (C.c_float * len(L.RT_STAGE_NAMES))()
        L.check(L.lib().bvh_build_times(self._h, ms))
        return dict(zip(L.RT_STAGE_NAMES, [float(x) for x in ms]))

    def device_buffers(self):
        """[(ptr, bytes)] of nodes, reordered triangles, header (for a broadcast)."""
        ptrs = (C.c_void_p * 3)()
        sizes = (C.c_uint64 * 3)()
        L.check(L.lib().bvh_device_buffers(self._h, ptrs, sizes))
        return [(int(ptrs[i] or 0), int(sizes[i])) for i in range(3)]

    def commit(self) -> None:
        L.check(L.lib().bvh_commit(self._h))

    # -- free_bvh ----------------------------------------------------------------
    def close(self) -> None:
        if self._h:
            L.lib().free_bvh(self._h)
            self._h = C.c_void_p()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


# ---------------------------------------------------------------------------
# single-process multi-GPU (build_bvh_multi / trace_rays_sharded / trace_random_sharded)
# ---------------------------------------------------------------------------
def shard_range(n: int, rank: int, world: int):
    """(first, count) of rank's share of n rays: the library's rule (rt_shard_range)."""
    first, count = C.c_uint64(), C.c_uint64()
    L.check(L.lib().rt_shard_range(n, rank, world, C.byref(first), C.byref(count)))
    return int(first.value), int(count.value)


class MultiBvh:
    """One BVH per entry of ``devices`` in this process; rays are sharded over them."""

    def __init__(self, handles):
        self.handles = handles            # list of Bvh
        self._arr = (C.c_void_p * len(handles))(*[h._h for h in handles])

    @classmethod
    def build(cls, tris, devices, key_bits: int = 30, strategy: int = L.RT_MULTI_REDUNDANT,
              flags: int = 0) -> "MultiBvh":
        """``flags``: RT_BUILD_* (the wide / compact layouts are derived on every device)."""
        n = _rows(tris, 12)
        opts = L.BuildOpts(key_bits=key_bits, flags=flags, device=-1, reserved=0, stream=None)
        devs = (C.c_int32 * len(devices))(*devices)
        out = (C.c_void_p * len(devices))()
        L.check(L.lib().build_bvh_multi(_ptr(tris) if n else None, n, C.byref(opts), devs, len(devices),
                                        strategy, out))
        return cls([Bvh(out[i]) for i in range(len(devices))])

    def trace(self, rays, flags: int = L.RT_TRACE_CLOSEST, out=None):
        """Host rays [n, 8] -> host hits [n, 4]; share i of the batch on handle i."""
        n = _rows(rays, 8)
        if out is None:
            out = _empty_like_kind(rays, (n, 4), None)
        L.check(L.lib().trace_rays_sharded(self._arr, len(self.handles), _ptr(rays), n, flags, _ptr(out)))
        return out

    def trace_random(self, box, seed: int, first: int, n_rays: int, flags: int = L.RT_TRACE_CLOSEST,
                     batch: int = 0, out=None):
        """Config 5: rays generated on the devices.  Returns (hits or None, n_hit[], trace_ms[])."""
        box = np.ascontiguousarray(box, dtype=np.float32).reshape(6)
        k = len(self.handles)
        n_hit, ms = (C.c_uint64 * k)(), (C.c_float * k)()
        L.check(L.lib().trace_random_sharded(self._arr, k, box.ctypes.data, seed, first, n_rays, flags,
                                             batch, _ptr(out) if out is not None else None, n_hit, ms))
        return out, [int(x) for x in n_hit], [float(x) for x in ms]

    def close(self) -> None:
        for h in self.handles:
            h.close()
        self.handles = []

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()


def merge_hits(parts, prim_base=None, flags: int = L.RT_TRACE_CLOSEST, out=None, stream: int = 0):
    """rt_merge_hits: per ray the best of the parts' hits by (t, prim + base); parts: CUDA tensors
    [n, 4] of the same rays traced against different parts of a scene."""
    k = len(parts)
    n = _rows(parts[0], 4)
    ptrs = (C.c_void_p * k)(*[p.data_ptr() for p in parts])
    base = (C.c_uint32 * k)(*([int(b) for b in prim_base] if prim_base is not None else [0] * k))
    if out is None:
        out = torch.empty((n, 4), dtype=torch.float32, device=parts[0].device)
    L.check(L.lib().rt_merge_hits(ptrs, base, k, n, flags, _ptr(out), stream or None))
    return out


class ShardedBvh:
    """The SCENE sharded over devices in this process (build_bvh_sharded / trace_rays_merged):
    part i holds a contiguous range of the triangles; hits are merged by (t, prim)."""

    def __init__(self, handles, prim_base):
        self.handles, self.prim_base = handles, list(prim_base)
        self._arr = (C.c_void_p * len(handles))(*[h._h for h in handles])
        self._base = (C.c_uint32 * len(handles))(*self.prim_base)

    @classmethod
    def build(cls, tris, devices, key_bits: int = 30) -> "ShardedBvh":
        n = _rows(tris, 12)
        opts = L.BuildOpts(key_bits=key_bits, flags=0, device=-1, reserved=0, stream=None)
        devs = (C.c_int32 * len(devices))(*devices)
        out = (C.c_void_p * len(devices))()
        base = (C.c_uint32 * len(devices))()
        L.check(L.lib().build_bvh_sharded(_ptr(tris) if n else None, n, C.byref(opts), devs, len(devices), out, base))
        return cls([Bvh(out[i]) for i in range(len(devices))], [int(b) for b in base])

    def trace(self, rays, flags: int = L.RT_TRACE_CLOSEST, out=None):
        n = _rows(rays, 8)
        if out is None:
            out = _empty_like_kind(rays, (n, 4), None)
        L.check(L.lib().trace_rays_merged(self._arr, self._base, len(self.handles), _ptr(rays), n, flags, _ptr(out)))
        return out

    def close(self) -> None:
        for h in self.handles:
            h.close()
        self.handles = []

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()


def trim(device: int = -1) -> None:
    """Returns the library's cached pool memory to the driver (rt_trim)."""
    L.check(L.lib().rt_trim(device))


def hit_prims(hits):
    """int32 view of the prim column."""
    if _is_tensor(hits):
        return hits[:, 3].contiguous().view(torch.int32)
    return np.ascontiguousarray(hits[:, 3]).view(np.int32)


def sort_pairs(keys, vals, key_bits: int = 30) -> float:
    """In-place device radix sort of (keys, vals); returns the device ms."""
    ms = C.c_float()
    n = int(keys.numel())
    L.check(L.lib().rt_sort_pairs(keys.data_ptr(), vals.data_ptr(), n, key_bits, C.byref(ms)))
    return float(ms.value)


def sort_pairs_static(keys, vals, stream: int = 0) -> None:
    """rt_sort_pairs_static: in-place device sort of (30-bit u32 keys, u32 payloads) that
    allocates nothing (static device scratch, at most RT_SORT_STATIC_MAX_PAIRS pairs);
    enqueued on ``stream``, not synchronised."""
    L.check(L.lib().rt_sort_pairs_static(keys.data_ptr(), vals.data_ptr(), int(keys.numel()), stream or None))


def set_param(name: str, value: int) -> None:
    L.check(L.lib().rt_set_param(name.encode(), value))


def kernel_launches() -> int:
    return int(L.lib().rt_kernel_launches())
