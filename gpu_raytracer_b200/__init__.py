"""gpu_raytracer_b200 — a B200-native LBVH ray-tracing core.

Hand-written sm_100a CUDA kernels (Morton encode, chunked LSD radix sort, fused
bottom-up Karras tree build + refit, optional BVH4 collapse, persistent while-while
traversal) behind the C ABI of
include/gpu_raytracer.h: build_bvh / trace_rays / free_bvh.  This package is the
ctypes mirror of that ABI plus host logic (camera, sharding, image output).
"""
from ._lib import (RT_ANY_HIT_PRIM, RT_SCENE_CORNELL, RT_SCENE_SOUP, RT_SCENE_SPHEREFLAKE,
                   RT_SCENE_TERRAIN, RT_TRACE_ANY, RT_TRACE_CLOSEST, RT_TRACE_REORDER, RT_TRACE_WATERTIGHT, RtError,
                   SO_PATH, lib)
from ._lib import RT_MULTI_BROADCAST, RT_MULTI_REDUNDANT
from ._lib import RT_BUILD_COMPACT, RT_BUILD_KEEP_KEYS, RT_BUILD_TIMINGS, RT_BUILD_WIDE4, RT_BUILD_WIDE8
from .api import (merge_hits, ShardedBvh, sort_pairs_static, MultiBvh, shard_range, trim, Bvh, gen_ao_rays, gen_primary_rays, gen_random_rays, gen_scene, hit_prims,
                  kernel_launches, load_obj, load_tris, save_tris, scene_size, set_param, shade,
                  sort_pairs)
from . import camera, image, sharding

__all__ = [
    "Bvh", "ShardedBvh", "merge_hits", "MultiBvh", "shard_range", "trim", "RT_MULTI_REDUNDANT", "RT_MULTI_BROADCAST", "gen_scene", "gen_primary_rays", "gen_random_rays", "gen_ao_rays", "shade", "scene_size",
    "hit_prims", "sort_pairs", "sort_pairs_static", "set_param", "load_obj", "load_tris", "save_tris", "kernel_launches", "camera", "image", "sharding", "lib",
    "RtError", "SO_PATH", "RT_TRACE_CLOSEST", "RT_TRACE_ANY", "RT_TRACE_WATERTIGHT", "RT_TRACE_REORDER", "RT_ANY_HIT_PRIM", "RT_SCENE_CORNELL",
    "RT_SCENE_SOUP", "RT_SCENE_SPHEREFLAKE", "RT_SCENE_TERRAIN",
    "RT_BUILD_KEEP_KEYS", "RT_BUILD_TIMINGS", "RT_BUILD_WIDE4", "RT_BUILD_WIDE8", "RT_BUILD_COMPACT",
]
