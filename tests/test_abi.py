"""CPU tests of the drop-in boundary: the shared library loads without a GPU,
exports every symbol include/gpu_raytracer.h declares, and refuses to compute
without a device (no CPU fallback)."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "gpu_raytracer.h")).read()
    return sorted(set(re.findall(r"^RT_API [\w\s\*]+?\b(\w+)\(", text, flags=re.M)))


def test_header_symbols_are_exported(rt):
    names = _declared_symbols()
    assert {"build_bvh", "trace_rays", "free_bvh"} <= set(names) and len(names) >= 20
    lib = ctypes.CDLL(rt.SO_PATH)
    for n in names:
        assert hasattr(lib, n), f"{n} is declared in the header but not exported"
    from gpu_raytracer_b200 import _lib
    assert set(_lib.SIGNATURES) == set(names), "ctypes table and header disagree"


def test_struct_layouts_match_the_header(rt):
    from gpu_raytracer_b200 import _lib
    assert ctypes.sizeof(_lib.BuildOpts) == 24
    assert ctypes.sizeof(_lib.BvhInfo) == 56
    assert _lib.BvhInfo.bounds.offset == 24 and _lib.BvhInfo.device_bytes.offset == 48


def test_scene_sizes_need_no_device(rt):
    assert rt.scene_size(rt.RT_SCENE_CORNELL) == 1024
    assert rt.scene_size(rt.RT_SCENE_SPHEREFLAKE, 4, 13) == 9979112
    assert rt.scene_size(rt.RT_SCENE_TERRAIN, 7072) == 100026368
    assert rt.lib().rt_version().startswith(b"gpu_raytracer_b200")


def test_argument_errors(rt):
    from gpu_raytracer_b200 import _lib
    L = rt.lib()
    h = ctypes.c_void_p()
    assert L.build_bvh(None, 4, None, ctypes.byref(h)) == _lib.RT_ERR_INVALID
    assert b"NULL" in L.bvh_last_error()
    opts = _lib.BuildOpts(key_bits=31, flags=0, device=-1, reserved=0, stream=None)
    t = np.zeros((4, 12), dtype=np.float32)
    assert L.build_bvh(t.ctypes.data, 4, ctypes.byref(opts), ctypes.byref(h)) == _lib.RT_ERR_INVALID
    assert L.build_bvh(t.ctypes.data, 1 << 30, None, ctypes.byref(h)) == _lib.RT_ERR_LIMIT
    assert L.trace_rays(None, t.ctypes.data, 1, 0, t.ctypes.data) == _lib.RT_ERR_INVALID
    assert L.rt_set_param(b"no.such.knob", 1) == _lib.RT_ERR_INVALID
    L.free_bvh(None)  # no-op


def test_no_cpu_fallback(rt):
    """Without a CUDA device every compute entry point fails loudly."""
    if rt.lib().rt_device_count() > 0:
        pytest.skip("a CUDA device is present")
    from gpu_raytracer_b200 import _lib
    with pytest.raises(rt.RtError) as e:
        rt.Bvh.build(np.zeros((4, 12), dtype=np.float32))
    assert e.value.code == _lib.RT_ERR_NO_DEVICE
    with pytest.raises(rt.RtError):
        rt.gen_scene(rt.RT_SCENE_CORNELL, device=None)


def test_product_does_not_import_the_oracle():
    """oracle/ is test infrastructure: nothing under gpu_raytracer_b200/ may use it."""
    pkg = os.path.join(ROOT, "gpu_raytracer_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                text = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(import|from)\s+oracle\b", text, flags=re.M), f
                assert "liblbvh_oracle" not in text and "oracle.cpp" not in text, f


def test_camera_and_shard_ranges(rt):
    cam = rt.camera.cornell_camera(256, 256)
    assert cam.shape == (12,) and cam.dtype == np.float32
    assert np.allclose(cam[:3], [0.5, 0.5, -1.5])
    for n in [0, 1, 31, 32, 33, 1000, 2073600, 10**9 + 7]:
        for world in [1, 2, 3, 4, 8]:
            pos = 0
            sizes = []
            for r in range(world):
                first, count = rt.sharding.shard_range(n, r, world)
                assert first == pos
                assert first % 32 == 0 or count == 0
                pos += count
                sizes.append(count)
            assert pos == n
            assert max(sizes) - min(sizes) < 64
