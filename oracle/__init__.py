"""ctypes binding of the spec-derived CPU oracle (oracle/oracle.cpp).

TEST INFRASTRUCTURE ONLY.  Import this from ``tests/``, from
``__graft_entry__.smoke()`` and from the ``cpu_baseline`` / ``--impl reference``
legs of ``bench.py`` — never from ``gpu_raytracer_b200``.

PARITY UNPINNED: /root/reference holds no code and no golden vectors
(README.md:1-2, test.txt:1), so this oracle is a scalar C++ statement of the
algorithm in BASELINE.json:5 / SURVEY.md §8(a), pinned by a brute-force
intersector, Morton known-answer vectors and std::stable_sort (see oracle.cpp).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liblbvh_oracle.so")

SCENE_CORNELL, SCENE_SOUP, SCENE_SPHEREFLAKE, SCENE_TERRAIN = 0, 1, 2, 3
TRACE_CLOSEST, TRACE_ANY, TRACE_WATERTIGHT = 0, 1, 2
ANY_HIT_PRIM = 0x7FFFFFFF


def build(force: bool = False) -> str:
    """Compile oracle.cpp with the committed Makefile; returns the .so path."""
    src = os.path.join(_HERE, "oracle.cpp")
    stale = (not os.path.exists(_SO)) or os.path.getmtime(_SO) < os.path.getmtime(src)
    if force or stale:
        env = dict(os.environ)
        try:
            with open("/proc/cpuinfo") as f:
                if " fma" not in f.read():
                    env["NOFMA"] = "1"
        except OSError:
            pass
        subprocess.run(["make", "-C", _HERE, "-B", "liblbvh_oracle.so"], check=True, env=env,
                       stdout=subprocess.DEVNULL)
    return _SO


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_SO)
        u64, u32, i32, f32p, vp = C.c_uint64, C.c_uint32, C.c_int, C.POINTER(C.c_float), C.c_void_p
        L.orc_set_threads.argtypes = [i32]
        L.orc_max_threads.restype = i32
        L.orc_morton30.argtypes = [u32, u32, u32]
        L.orc_morton30.restype = u32
        L.orc_morton63.argtypes = [u32, u32, u32]
        L.orc_morton63.restype = u64
        L.orc_scene_size.argtypes = [i32, u64, u64]
        L.orc_scene_size.restype = u64
        L.orc_gen_scene.argtypes = [i32, u64, u64, u64, u64, u64, vp]
        L.orc_gen_scene.restype = i32
        L.orc_gen_primary_rays.argtypes = [vp, u32, u32, u64, u64, vp]
        L.orc_gen_random_rays.argtypes = [vp, u64, u64, u64, vp]
        L.orc_gen_ao_rays.argtypes = [vp, vp, vp, u64, u32, u64, C.c_float, vp]
        L.orc_shade.argtypes = [vp, vp, vp, vp, u64, u32, vp]
        L.orc_build.argtypes = [vp, u64, u32]
        L.orc_build.restype = vp
        L.orc_build_borrow.argtypes = [vp, u64, u32]
        L.orc_build_borrow.restype = vp
        L.orc_free.argtypes = [vp]
        L.orc_read.argtypes = [vp] * 8
        L.orc_build_times.argtypes = [vp, vp]
        L.orc_trace.argtypes = [vp, vp, u64, u32, vp, vp]
        L.orc_brute.argtypes = [vp, u64, vp, u64, u32, vp]
        L.orc_collapse4.argtypes = [vp, u32, u64, vp, vp]
        L.orc_collapse4.restype = u64
        L.orc_collapse8.argtypes = [vp, u64, vp, vp]
        L.orc_collapse8.restype = u64
        L.orc_canon8.argtypes = [vp, u64, vp, u64, vp, vp]
        L.orc_canon8.restype = u64
        _lib = L
    return _lib


def _p(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


def set_threads(n: int) -> None:
    lib().orc_set_threads(int(n))


def max_threads() -> int:
    return int(lib().orc_max_threads())


def morton30(x, y, z) -> int:
    return int(lib().orc_morton30(x, y, z))


def morton63(x, y, z) -> int:
    return int(lib().orc_morton63(x, y, z))


def scene_size(kind: int, p0: int = 0, p1: int = 0) -> int:
    return int(lib().orc_scene_size(kind, p0, p1))


def gen_scene(kind: int, p0: int = 0, p1: int = 0, seed: int = 0, first: int = 0,
              count: int | None = None) -> np.ndarray:
    """Triangles [first, first+count) as float32 [count, 12]."""
    if count is None:
        count = scene_size(kind, p0, p1) - first
    out = np.empty((count, 12), dtype=np.float32)
    rc = lib().orc_gen_scene(kind, p0, p1, seed, first, count, _p(out))
    if rc != 0:
        raise ValueError(f"orc_gen_scene failed ({rc})")
    return out


def gen_primary_rays(cam: np.ndarray, width: int, height: int, first: int = 0,
                     count: int | None = None) -> np.ndarray:
    cam = np.ascontiguousarray(cam, dtype=np.float32).reshape(12)
    if count is None:
        count = width * height - first
    out = np.empty((count, 8), dtype=np.float32)
    lib().orc_gen_primary_rays(_p(cam), width, height, first, count, _p(out))
    return out


def gen_random_rays(box: np.ndarray, seed: int, first: int, count: int) -> np.ndarray:
    box = np.ascontiguousarray(box, dtype=np.float32).reshape(6)
    out = np.empty((count, 8), dtype=np.float32)
    lib().orc_gen_random_rays(_p(box), seed, first, count, _p(out))
    return out


def gen_ao_rays(tris: np.ndarray, rays: np.ndarray, hits: np.ndarray, spp: int, seed: int,
                radius: float) -> np.ndarray:
    n = rays.shape[0]
    out = np.empty((n * spp, 8), dtype=np.float32)
    lib().orc_gen_ao_rays(_p(tris), _p(rays), _p(hits), n, spp, seed, C.c_float(radius), _p(out))
    return out


def shade(tris: np.ndarray, rays: np.ndarray, hits: np.ndarray, ao_hits=None, spp: int = 0) -> np.ndarray:
    """§8(f).1: grey = |n.d| * (1 - occluded/spp), float32 [n, 3]; tris in original order."""
    tris = np.ascontiguousarray(tris, dtype=np.float32)
    rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 8)
    hits = np.ascontiguousarray(hits, dtype=np.float32).reshape(-1, 4)
    n = rays.shape[0]
    out = np.empty((n, 3), dtype=np.float32)
    if ao_hits is not None and spp:
        ao_hits = np.ascontiguousarray(ao_hits, dtype=np.float32).reshape(-1, 4)
        assert ao_hits.shape[0] == n * spp
    lib().orc_shade(_p(tris), _p(rays), _p(hits), _p(ao_hits) if ao_hits is not None and spp else None,
                    n, spp if ao_hits is not None else 0, _p(out))
    return out


def compact_grid(bounds: np.ndarray) -> np.ndarray:
    """Grid of the 32-byte binary layout (DESIGN.md §5 "compact nodes"; no reference counterpart):
    per axis 65,000 steps across the scene box, never finer than 2^-21 of the box's largest
    coordinate, the step rounded up to float; plane 0 lies three steps below the box.
    Returns [origin xyz, step xyz] as float32."""
    b = np.asarray(bounds, dtype=np.float64)
    lo, hi = b[:3], b[3:]
    step = np.maximum(np.maximum((hi - lo) / 65000.0, np.maximum(np.abs(lo), np.abs(hi)) * 2.0 ** -21), 1e-30)
    sf = step.astype(np.float32)
    up = sf.astype(np.float64) < step
    sf[up] = np.nextafter(sf[up], np.float32(np.inf))
    origin = (lo - 3.0 * sf.astype(np.float64)).astype(np.float32)
    return np.concatenate([origin, sf]).astype(np.float32)


def compact_pack(boxes: np.ndarray, grid: np.ndarray) -> np.ndarray:
    """16-bit plane pairs of [m, 6] float32 boxes on ``grid``: per axis lo | hi << 16, both rounded
    outwards by at least one whole step (floor(x - 1e-6) - 1, ceil(x + 1e-6) + 1 in float64, x in
    steps from plane 0).  Returns uint32 [m, 3]."""
    boxes = np.asarray(boxes, dtype=np.float32)
    origin = grid[:3].astype(np.float64)
    inv = 1.0 / grid[3:].astype(np.float64)
    xl = (boxes[:, :3].astype(np.float64) - origin) * inv
    xh = (boxes[:, 3:].astype(np.float64) - origin) * inv
    ql = np.clip(np.floor(xl - 1e-6).astype(np.int64) - 1, 0, 65535)
    qh = np.clip(np.ceil(xh + 1e-6).astype(np.int64) + 1, 0, 65535)
    return (ql | (qh << 16)).astype(np.uint32)


def canon8(nodes: np.ndarray, tris: np.ndarray):
    """A compressed 8-wide tree (bvh_read_wide8) renumbered breadth first, the numbering
    ``OracleBvh.collapse8`` uses.  Returns (nodes uint8 [m, 80], tris float32 [n, 12])."""
    nodes = np.ascontiguousarray(nodes).view(np.uint8).reshape(-1, 80)
    tris = np.ascontiguousarray(tris, dtype=np.float32).reshape(-1, 12)
    out_n, out_t = np.zeros_like(nodes), np.zeros_like(tris)
    m = int(lib().orc_canon8(_p(nodes), nodes.shape[0], _p(tris), tris.shape[0], _p(out_n), _p(out_t)))
    if m != nodes.shape[0]:
        raise ValueError(f"malformed 8-wide tree: {m} of {nodes.shape[0]} nodes reached")
    return out_n, out_t


class OracleBvh:
    """Result of the oracle's build: canonical arrays."""

    def __init__(self, tris: np.ndarray, key_bits: int = 30, lazy: bool = False):
        """lazy=True (the 100M-triangle test): the oracle borrows ``tris`` instead of copying
        it, and the result arrays are fetched one at a time with ``read(name)``."""
        tris = np.ascontiguousarray(tris, dtype=np.float32).reshape(-1, 12)
        self.n = int(tris.shape[0])
        self.key_bits = key_bits
        self._tris = tris
        self._h = (lib().orc_build_borrow if lazy else lib().orc_build)(_p(tris), self.n, key_bits)
        if not self._h:
            raise ValueError("orc_build failed")
        t = np.zeros(4, dtype=np.float64)
        lib().orc_build_times(self._h, _p(t))
        self.times = dict(morton=t[0], sort=t[1], topology=t[2], refit=t[3], total=float(t.sum()))
        if lazy:
            return
        for name in self._ARRAYS:
            setattr(self, name, self.read(name))

    _ARRAYS = ("keys", "idx", "left", "right", "parent", "boxes", "bounds")

    def read(self, name: str) -> np.ndarray:
        """One result array of the build (a fresh numpy copy)."""
        n = self.n
        nodes = max(2 * n - 1, 0)
        shape, dtype = {"keys": ((n,), np.uint64), "idx": ((n,), np.uint32),
                        "left": ((max(n - 1, 0),), np.uint32), "right": ((max(n - 1, 0),), np.uint32),
                        "parent": ((nodes,), np.uint32), "boxes": ((nodes, 6), np.float32),
                        "bounds": ((6,), np.float32)}[name]
        out = np.zeros(shape, dtype=dtype)
        args = [None] * 7
        args[self._ARRAYS.index(name)] = _p(out)
        lib().orc_read(self._h, *args)
        return out

    def trace(self, rays: np.ndarray, flags: int = TRACE_CLOSEST, counters: bool = False):
        rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 8)
        hits = np.empty((rays.shape[0], 4), dtype=np.float32)
        cnt = np.zeros(2, dtype=np.uint64)
        lib().orc_trace(self._h, _p(rays), rays.shape[0], flags, _p(hits), _p(cnt))
        if counters:
            return hits, (int(cnt[0]), int(cnt[1]))
        return hits

    def collapse4(self, leaf_max: int = 4):
        """A5w: the 4-wide tree in pre-order: (boxes [m,4,6] f32, desc [m,4,3] u32)."""
        m = int(lib().orc_collapse4(self._h, leaf_max, 0, None, None))
        boxes = np.zeros((m, 4, 6), dtype=np.float32)
        desc = np.zeros((m, 4, 3), dtype=np.uint32)
        if m:
            lib().orc_collapse4(self._h, leaf_max, m, _p(boxes), _p(desc))
        return boxes, desc

    def collapse8(self):
        """A5w: the compressed 8-wide layout: (nodes uint8 [m, 80], tris float32 [n, 12])."""
        m = int(lib().orc_collapse8(self._h, 0, None, None))
        nodes = np.zeros((m, 80), dtype=np.uint8)
        tris = np.zeros((self.n, 12), dtype=np.float32)
        if m:
            lib().orc_collapse8(self._h, m, _p(nodes), _p(tris))
        return nodes, tris

    def close(self):
        if self._h:
            lib().orc_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def brute(tris: np.ndarray, rays: np.ndarray, flags: int = TRACE_CLOSEST) -> np.ndarray:
    tris = np.ascontiguousarray(tris, dtype=np.float32).reshape(-1, 12)
    rays = np.ascontiguousarray(rays, dtype=np.float32).reshape(-1, 8)
    hits = np.empty((rays.shape[0], 4), dtype=np.float32)
    lib().orc_brute(_p(tris), tris.shape[0], _p(rays), rays.shape[0], flags, _p(hits))
    return hits


def hit_prims(hits: np.ndarray) -> np.ndarray:
    return np.ascontiguousarray(hits[:, 3]).view(np.int32)
