"""ctypes loader for libgpu_raytracer_b200.so (include/gpu_raytracer.h).

The library is the product; this module only declares its C ABI.  If the shared
object is missing it raises — there is no Python or CPU fallback.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# RT_LIB: development override used to A/B experimental builds of the same library
SO_PATH = os.environ.get("RT_LIB") or os.path.join(_HERE, "libgpu_raytracer_b200.so")

RT_OK, RT_ERR_INVALID, RT_ERR_CUDA, RT_ERR_OOM, RT_ERR_NO_DEVICE, RT_ERR_LIMIT = 0, -1, -2, -3, -4, -5
RT_TRACE_CLOSEST, RT_TRACE_ANY, RT_TRACE_WATERTIGHT, RT_TRACE_REORDER = 0, 1, 2, 4
RT_ANY_HIT_PRIM = 0x7FFFFFFF
RT_BUILD_KEEP_KEYS, RT_BUILD_TIMINGS, RT_BUILD_WIDE4, RT_BUILD_WIDE8, RT_BUILD_COMPACT = 1, 2, 4, 8, 16
RT_SCENE_CORNELL, RT_SCENE_SOUP, RT_SCENE_SPHEREFLAKE, RT_SCENE_TERRAIN = 0, 1, 2, 3
RT_MULTI_REDUNDANT, RT_MULTI_BROADCAST = 0, 1
RT_STAGE_NAMES = ("bounds", "morton", "sort", "topology", "refit", "total", "wide")
RT_MAX_TRIS = (1 << 30) - 1
RT_SORT_STATIC_MAX_PAIRS = 1 << 20


class BuildOpts(C.Structure):
    _fields_ = [("key_bits", C.c_uint32), ("flags", C.c_uint32), ("device", C.c_int32),
                ("reserved", C.c_int32), ("stream", C.c_void_p)]


class BvhInfo(C.Structure):
    _fields_ = [("n_tris", C.c_uint64), ("n_nodes", C.c_uint64), ("key_bits", C.c_uint32),
                ("device", C.c_int32), ("bounds", C.c_float * 6), ("device_bytes", C.c_uint64)]


# name -> (restype, argtypes); every symbol include/gpu_raytracer.h declares
_u64, _u32, _i32, _vp, _f = C.c_uint64, C.c_uint32, C.c_int, C.c_void_p, C.c_float
SIGNATURES = {
    "build_bvh": (_i32, [_vp, _u64, C.POINTER(BuildOpts), C.POINTER(_vp)]),
    "trace_rays": (_i32, [_vp, _vp, _u64, _u32, _vp]),
    "free_bvh": (None, [_vp]),
    "bvh_last_error": (C.c_char_p, []),
    "trace_rays_async": (_i32, [_vp, _vp, _u64, _u32, _vp, _vp]),
    "trace_rays_stats": (_i32, [_vp, _vp, _u64, _u32, _vp, _vp]),
    "bvh_info": (_i32, [_vp, C.POINTER(BvhInfo)]),
    "bvh_read_sorted": (_i32, [_vp, _vp, _vp]),
    "bvh_read_topology": (_i32, [_vp, _vp, _vp, _vp]),
    "bvh_read_boxes": (_i32, [_vp, _vp]),
    "bvh_build_times": (_i32, [_vp, _vp]),
    "refit_bvh": (_i32, [_vp, _vp]),
    "bvh_widen": (_i32, [_vp]),
    "bvh_read_wide": (_i32, [_vp, _vp, C.POINTER(_u32)]),
    "bvh_widen8": (_i32, [_vp]),
    "bvh_compact": (_i32, [_vp]),
    "bvh_read_compact": (_i32, [_vp, _vp, _vp]),
    "bvh_read_wide8": (_i32, [_vp, _vp, _vp, C.POINTER(_u32)]),
    "bvh_device_buffers": (_i32, [_vp, C.POINTER(_vp), C.POINTER(_u64)]),
    "bvh_alloc_empty": (_i32, [_u64, _u32, C.POINTER(_vp)]),
    "bvh_commit": (_i32, [_vp]),
    "rt_load_obj": (_i32, [C.c_char_p, C.POINTER(_vp), C.POINTER(_u64)]),
    "rt_save_tris": (_i32, [C.c_char_p, _vp, _u64]),
    "rt_load_tris": (_i32, [C.c_char_p, C.POINTER(_vp), C.POINTER(_u64)]),
    "rt_free_host": (None, [_vp]),
    "bvh_save": (_i32, [_vp, C.c_char_p]),
    "bvh_load": (_i32, [C.c_char_p, C.POINTER(_vp)]),
    "rt_scene_size": (_u64, [_i32, _u64, _u64]),
    "rt_gen_scene": (_i32, [_i32, _u64, _u64, _u64, _u64, _u64, _vp]),
    "rt_gen_primary_rays": (_i32, [_vp, _u32, _u32, _u64, _u64, _vp]),
    "rt_gen_random_rays": (_i32, [_vp, _u64, _u64, _u64, _vp]),
    "rt_gen_ao_rays": (_i32, [_vp, _vp, _vp, _u64, _u32, _u64, _f, _vp]),
    "rt_shade": (_i32, [_vp, _vp, _vp, _vp, _u64, _u32, _vp]),
    "rt_render_ao": (_i32, [_vp, _vp, _vp, _u64, _u32, _u64, _f, _u32, _vp, _vp]),
    "rt_render_camera_ao": (_i32, [_vp, _vp, _vp, _u32, _u32, _u32, _u64, _f, _u32, _vp, _vp]),
    "rt_device_count": (_i32, []),
    "rt_kernel_launches": (_u64, []),
    "rt_set_param": (_i32, [C.c_char_p, C.c_int64]),
    "rt_sort_pairs": (_i32, [_vp, _vp, _u64, _u32, _vp]),
    "rt_trim": (_i32, [_i32]),
    "rt_merge_hits": (_i32, [C.POINTER(_vp), C.POINTER(_u32), _i32, _u64, _u32, _vp, _vp]),
    "build_bvh_sharded": (_i32, [_vp, _u64, C.POINTER(BuildOpts), C.POINTER(_i32), _i32, C.POINTER(_vp),
                                 C.POINTER(_u32)]),
    "trace_rays_merged": (_i32, [C.POINTER(_vp), C.POINTER(_u32), _i32, _vp, _u64, _u32, _vp]),
    "rt_sort_pairs_static": (_i32, [_vp, _vp, _u64, _vp]),
    "trace_camera": (_i32, [_vp, _vp, _u32, _u32, _u64, _u64, _u32, _vp]),
    "rt_shard_range": (_i32, [_u64, _i32, _i32, C.POINTER(_u64), C.POINTER(_u64)]),
    "build_bvh_multi": (_i32, [_vp, _u64, C.POINTER(BuildOpts), C.POINTER(_i32), _i32, _i32,
                               C.POINTER(_vp)]),
    "trace_rays_sharded": (_i32, [C.POINTER(_vp), _i32, _vp, _u64, _u32, _vp]),
    "trace_random_sharded": (_i32, [C.POINTER(_vp), _i32, _vp, _u64, _u64, _u64, _u32, _u64, _vp,
                                    C.POINTER(_u64), C.POINTER(_f)]),
    "rt_version": (C.c_char_p, []),
}

_lib = None


class RtError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"gpu_raytracer_b200 error {code}: {msg}")
        self.code = code


def lib() -> C.CDLL:
    """Loads the CUDA library; raises if it has not been built (no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(SO_PATH):
            raise ImportError(
                f"{SO_PATH} is missing: build it with `python -m gpu_raytracer_b200.build` "
                "(or __graft_entry__.build()); gpu_raytracer_b200 has no CPU/PyTorch fallback")
        L = C.CDLL(SO_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)  # AttributeError if the ABI is incomplete
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(rc: int) -> None:
    if rc != RT_OK:
        raise RtError(rc, lib().bvh_last_error().decode("utf-8", "replace"))
