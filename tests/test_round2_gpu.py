"""GPU parity tests added in round 2 (VERDICT r1 "Next round" items 3, 8, 9 and ADVICE r1):

  config 4 at full size   100,026,368-triangle terrain, 63-bit (hybrid order) and 30-bit builds,
                          every array byte-equal to the oracle's
  config 5 samples        2^22-ray windows of the 10^9-ray sequence against the 10M flake, plain
                          and RT_TRACE_REORDER, bit-equal to the oracle
  T6 on one GPU           shard ranges traced separately == one call; the C-ABI multi-GPU entry
                          with device lists [0] and [0, 0], both strategies
  shading                 rt_shade bit-equal to the oracle's shade; the frame's PPM sha256
  zero direction          rays with +-0 components on faces / edges / vertices == brute force
  files                   corrupt and truncated BVH files are RT_ERR_INVALID, not a fault
"""
import ctypes
import gc
import hashlib
import json
import os
import subprocess
import threading

import numpy as np
import pytest

from conftest import bits
from test_host_logic import zero_direction_cases

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def _mem_available_gb():
    try:
        for line in open("/proc/meminfo"):
            if line.startswith("MemAvailable:"):
                return int(line.split()[1]) / 1e6
    except OSError:
        pass
    return 0.0


# ---- config 4 at full size ------------------------------------------------------------------
@pytest.mark.parametrize("key_bits", [63, 30])
def test_config4_hundred_million(rt, oracle, key_bits):
    """Config 4 (BASELINE.json:10): the 100,026,368-triangle terrain against the CPU oracle, array
    by array (the arrays are fetched and compared one at a time to bound host memory: ~25 GB)."""
    if _mem_available_gb() < 40:
        pytest.skip(f"needs ~25 GB of host memory, {_mem_available_gb():.0f} GB available")
    d_tris = rt.gen_scene(rt.RT_SCENE_TERRAIN, 7072, 0, 3)
    n = int(d_tris.shape[0])
    assert n == 100026368
    tris = d_tris.cpu().numpy()
    # the device generator against the oracle's on three windows (full equality: other tests)
    for first in (0, 50_000_000, n - 4096):
        assert np.array_equal(bits(tris[first:first + 4096]),
                              bits(oracle.gen_scene(oracle.SCENE_TERRAIN, 7072, 0, 3, first, 4096)))
    ob = oracle.OracleBvh(tris, key_bits, lazy=True)
    try:
        with rt.Bvh.build(d_tris, key_bits=key_bits, keep_keys=True) as gb:
            del d_tris
            assert np.array_equal(bits(gb.bounds), bits(ob.read("bounds"))), "A1 scene bounds"
            keys, idx = gb.read_sorted()
            assert (keys[1:] >= keys[:-1]).all(), "sortedness"
            assert np.array_equal(keys, ob.read("keys")), "A2/A3 sorted Morton codes"
            if key_bits == 63:
                dup = int((keys[1:] == keys[:-1]).sum())
                assert dup < n // 2, "63-bit keys are (mostly) distinct"
            del keys
            assert np.array_equal(idx, ob.read("idx")), "A3 sorted primitive indices"
            del idx
            left, right, parent = gb.read_topology()
            assert np.array_equal(left, ob.read("left")), "A4 left children"
            assert np.array_equal(right, ob.read("right")), "A4 right children"
            assert np.array_equal(parent, ob.read("parent")), "A4 parents"
            del left, right, parent
            gc.collect()
            boxes = gb.read_boxes()
            want = ob.read("boxes")
            assert np.array_equal(bits(boxes), bits(want)), "A5 node boxes"
            del boxes, want
    finally:
        ob.close()
        rt.trim()


# ---- config 5 samples -------------------------------------------------------------------------
def test_config5_sample(rt, oracle):
    """Config 5 (BASELINE.json:11): windows of 2^22 rays at the start, the middle and the end of
    the 10^9-ray sequence, against the 10M-triangle sphereflake: plain and RT_TRACE_REORDER
    launches give the oracle's bits."""
    from gpu_raytracer_b200.configs import WORKLOADS
    wl = WORKLOADS["random1b"]
    d_tris = rt.gen_scene(*wl.scene)
    tris = d_tris.cpu().numpy()
    ob = oracle.OracleBvh(tris, wl.key_bits)
    m = 1 << 22
    total_hits = 0
    with rt.Bvh.build(d_tris, key_bits=wl.key_bits) as gb:
        assert np.array_equal(bits(gb.bounds), bits(ob.bounds))
        for first in (0, 5 * 10**8, 10**9 - m):
            g_rays = rt.gen_random_rays(gb.bounds, wl.random_seed, first, m)
            rays = g_rays.cpu().numpy()
            assert np.array_equal(bits(rays[:4096]),
                                  bits(oracle.gen_random_rays(ob.bounds, wl.random_seed, first, 4096)))
            want = ob.trace(rays)
            plain = gb.trace(g_rays).cpu().numpy()
            assert np.array_equal(bits(plain), bits(want)), first
            reordered = gb.trace(g_rays, rt.RT_TRACE_REORDER).cpu().numpy()
            assert np.array_equal(bits(reordered), bits(want)), first
            total_hits += int((oracle.hit_prims(want) >= 0).sum())
    assert 0.05 < total_hits / (3 * m) < 0.95
    ob.close()


# ---- T6 on one GPU --------------------------------------------------------------------------------
def test_sharded_ranges_equal_one_call(rt, oracle):
    """T6 with the ranks emulated on one GPU: every rank's range of
    sharding.trace_sharded, traced separately and concatenated, equals one call on the whole
    batch and the oracle, for world sizes 1, 2, 3 and 8 (tools/mgpu_check.py's assertions)."""
    from gpu_raytracer_b200 import sharding
    n_rays = 1_000_003
    tris = rt.gen_scene(rt.RT_SCENE_SPHEREFLAKE, 3, 6, 0)
    ob = oracle.OracleBvh(tris.cpu().numpy())
    with rt.Bvh.build(tris) as bvh:
        box = bvh.bounds
        whole = bvh.trace(rt.gen_random_rays(box, 4, 0, n_rays))
        want = ob.trace(oracle.gen_random_rays(ob.bounds, 4, 0, n_rays))
        assert np.array_equal(bits(whole.cpu().numpy()), bits(want))
        for world in (1, 2, 3, 8):
            parts = []
            for rank in range(world):
                for first, c, h in sharding.trace_sharded(
                        bvh.trace, lambda f, c: rt.gen_random_rays(box, 4, f, c), n_rays, rank, world,
                        batch=300_000):
                    parts.append(h)
                # the same shard in one call with RT_TRACE_REORDER: same bytes
                f0, c0 = sharding.shard_range(n_rays, rank, world)
                again = bvh.trace(rt.gen_random_rays(box, 4, f0, c0), rt.RT_TRACE_REORDER)
                assert torch.equal(again.view(torch.int32), whole[f0:f0 + c0].view(torch.int32))
            full = torch.cat(parts)
            assert torch.equal(full.view(torch.int32), whole.view(torch.int32)), world
    ob.close()


@pytest.mark.parametrize("strategy", ["redundant", "broadcast"])
def test_multi_gpu_handles_keep_the_extra_layouts(rt, oracle, strategy):
    """RT_BUILD_WIDE4 / WIDE8 / COMPACT with a device list: every handle holds the layouts, also
    the ones filled by a peer copy (derived again on the receiving device), with the bytes of
    the first handle's."""
    devices = [0, torch.cuda.device_count() - 1, 0]
    strat = rt.RT_MULTI_REDUNDANT if strategy == "redundant" else rt.RT_MULTI_BROADCAST
    tris = oracle.gen_scene(oracle.SCENE_SPHEREFLAKE, 3, 4, 0)
    rays = oracle.gen_random_rays(oracle.OracleBvh(tris).bounds, 4, 0, 100_000)
    want = oracle.OracleBvh(tris).trace(rays)
    for flag, read in ((rt.RT_BUILD_COMPACT, "read_compact"), (rt.RT_BUILD_WIDE8, "read_wide8"), (rt.RT_BUILD_WIDE4, "read_wide")):
        with rt.MultiBvh.build(dev(tris), devices, strategy=strat, flags=flag) as mb:
            # wide8 allocates nodes in a varying order (canon8 renumbers); BVH4 leaves the slots of
            # small subtrees unwritten: there only the root reference is compared
            view = {"read_compact": lambda t: t, "read_wide8": lambda t: oracle.canon8(*t),
                    "read_wide": lambda t: (np.array([t[1]]),)}[read]
            first = view(getattr(mb.handles[0], read)())
            for h in mb.handles[1:]:
                got = view(getattr(h, read)())
                assert all(np.array_equal(np.asarray(a).view(np.uint8), np.asarray(b).view(np.uint8))
                           for a, b in zip(first, got)), read
            assert np.array_equal(bits(mb.trace(rays)), bits(want)), read


@pytest.mark.parametrize("devices", [[0], [0, 0], [0, 0, 0]])
@pytest.mark.parametrize("strategy", ["redundant", "broadcast"])
def test_multi_gpu_entry_of_the_c_abi(rt, oracle, devices, strategy):
    """build_bvh_multi / trace_rays_sharded / trace_random_sharded (SURVEY.md §8b "Multi-GPU entry
    takes a device list"), on the GPUs this box has: device lists that name GPU 0 one to three
    times exercise the sharding, the peer copies and the per-device threads; with more GPUs
    visible the last slot goes to the last GPU."""
    if torch.cuda.device_count() > 1:
        devices = devices[:-1] + [torch.cuda.device_count() - 1]
    strat = rt.RT_MULTI_REDUNDANT if strategy == "redundant" else rt.RT_MULTI_BROADCAST
    tris = oracle.gen_scene(oracle.SCENE_SPHEREFLAKE, 3, 4, 0)
    ob = oracle.OracleBvh(tris)
    n_rays = 700_001
    rays = oracle.gen_random_rays(ob.bounds, 4, 0, n_rays)
    want = ob.trace(rays)
    for src in (tris, dev(tris)):      # host triangles, device triangles
        with rt.MultiBvh.build(src, devices, strategy=strat) as mb:
            assert len(mb.handles) == len(devices)
            for h in mb.handles:
                assert h.n_tris == ob.n and np.array_equal(bits(h.bounds), bits(ob.bounds))
                left, right, parent = h.read_topology()
                assert np.array_equal(left, ob.left) and np.array_equal(parent, ob.parent)
            got = mb.trace(rays)
            assert np.array_equal(bits(got), bits(want))
            pinned = torch.from_numpy(rays).pin_memory()
            out = torch.empty((n_rays, 4), dtype=torch.float32).pin_memory()
            mb.trace(pinned, out=out)
            assert np.array_equal(bits(out.numpy()), bits(want))
            occ = mb.trace(rays, rt.RT_TRACE_ANY)
            assert np.array_equal(bits(occ), bits(ob.trace(rays, oracle.TRACE_ANY)))
            # generated on the devices, hits gathered on the host
            hits = np.zeros((n_rays, 4), dtype=np.float32)
            for flags in (0, rt.RT_TRACE_REORDER):
                hits[:] = 0
                _, n_hit, ms = mb.trace_random(ob.bounds, 4, 0, n_rays, flags, batch=100_000, out=hits)
                assert np.array_equal(bits(hits), bits(want)), flags
                assert sum(n_hit) == int((oracle.hit_prims(want) >= 0).sum())
                assert all(m > 0 for m in ms)
                for i in range(len(devices)):
                    f, c = rt.shard_range(n_rays, i, len(devices))
                    assert n_hit[i] == int((oracle.hit_prims(want[f:f + c]) >= 0).sum())
            # a window in the middle of the sequence, hits left on the device
            _, n_hit, _ = mb.trace_random(ob.bounds, 4, 123_457, 50_000, 0)
            w = ob.trace(oracle.gen_random_rays(ob.bounds, 4, 123_457, 50_000))
            assert sum(n_hit) == int((oracle.hit_prims(w) >= 0).sum())
    ob.close()
    with pytest.raises(rt.RtError):
        rt.MultiBvh.build(tris, [0, 99])
    with pytest.raises(rt.RtError):
        rt.MultiBvh.build(tris, [])


# ---- trace_camera ------------------------------------------------------------------------------------
def test_trace_camera_equals_generated_rays(rt, oracle):
    tris = oracle.gen_scene(oracle.SCENE_CORNELL)
    cam = rt.camera.cornell_camera(640, 480)
    ob = oracle.OracleBvh(tris)
    want = ob.trace(oracle.gen_primary_rays(cam, 640, 480))
    with rt.Bvh.build(dev(tris)) as gb:
        got = gb.trace_camera(cam, 640, 480)                                 # numpy out
        assert isinstance(got, np.ndarray) and np.array_equal(bits(got), bits(want))
        got = gb.trace_camera(cam, 640, 480, device="cuda")                  # device out
        assert np.array_equal(bits(got.cpu().numpy()), bits(want))
        out = torch.empty((1000, 4), dtype=torch.float32).pin_memory()       # pinned, a sub-range
        gb.trace_camera(cam, 640, 480, out=out, first=12345, count=1000)
        assert np.array_equal(bits(out.numpy()), bits(want[12345:13345]))
        occ = gb.trace_camera(cam, 640, 480, rt.RT_TRACE_ANY)
        assert np.array_equal(oracle.hit_prims(occ) >= 0, oracle.hit_prims(want) >= 0)
        with pytest.raises(rt.RtError):
            gb.trace_camera(cam, 640, 480, first=640 * 480, count=1)
    ob.close()


def test_render_ao_equals_the_separate_calls(rt):
    """rt_render_ao (banded, three streams) gives the bytes of trace_rays -> rt_gen_ao_rays ->
    trace_rays(any) -> rt_shade, for device, pageable and pinned buffers, over several bands
    (1.44M rays > 1M per band) and with a ragged last band."""
    w, h, spp, seed = 1600, 901, 3, 77
    cam = rt.camera.cornell_camera(w, h)
    d_tris = rt.gen_scene(rt.RT_SCENE_CORNELL)
    for flags in (0, rt.RT_TRACE_WATERTIGHT):
        with rt.Bvh.build(d_tris) as gb:
            rays = rt.gen_primary_rays(cam, w, h)
            radius = float(np.float32(0.05) * np.float32(np.linalg.norm(gb.bounds[3:] - gb.bounds[:3])))
            hits = gb.trace(rays, flags)
            ao = rt.gen_ao_rays(d_tris, rays, hits, spp, seed, radius)
            occ = gb.trace(ao, flags | rt.RT_TRACE_ANY)
            rgb = rt.shade(d_tris, rays, hits, occ, spp)
            want_h, want_g = bits(hits.cpu().numpy()), bits(rgb.cpu().numpy())
            got_h, got_g = gb.render_ao(d_tris, rays, spp, seed, radius, flags)                 # device
            assert got_h.is_cuda and np.array_equal(bits(got_h.cpu().numpy()), want_h)
            assert np.array_equal(bits(got_g.cpu().numpy()), want_g)
            got_h, got_g = gb.render_ao(d_tris, rays.cpu().numpy(), spp, seed, radius, flags)   # pageable
            assert isinstance(got_g, np.ndarray)
            assert np.array_equal(bits(got_h), want_h) and np.array_equal(bits(got_g), want_g)
            pinned = rays.cpu().pin_memory()
            for _ in range(2):                                                                  # slots are reused
                got_h, got_g = gb.render_ao(d_tris, pinned, spp, seed, radius, flags)
                assert got_g.is_pinned()
                assert np.array_equal(bits(got_h.numpy()), want_h) and np.array_equal(bits(got_g.numpy()), want_g)
            # mixed: host rays, device frame; and no occlusion pass at all
            d_rgb = torch.empty((w * h, 3), dtype=torch.float32, device="cuda")
            gb.render_ao(d_tris, pinned, spp, seed, radius, flags, rgb=d_rgb)
            assert np.array_equal(bits(d_rgb.cpu().numpy()), want_g)
            _, plain = gb.render_ao(d_tris, pinned, 0, seed, radius, flags)
            assert np.array_equal(bits(plain.numpy()), bits(rt.shade(d_tris, rays, hits).cpu().numpy()))
            # the same frame from the camera alone: numpy out, pinned out, device out
            ch, cg = gb.render_camera_ao(d_tris, cam, w, h, spp, seed, radius, flags)
            assert isinstance(cg, np.ndarray) and np.array_equal(bits(ch), want_h) and np.array_equal(bits(cg), want_g)
            ph = torch.empty((w * h, 4), dtype=torch.float32).pin_memory()
            pg = torch.empty((w * h, 3), dtype=torch.float32).pin_memory()
            gb.render_camera_ao(d_tris, cam, w, h, spp, seed, radius, flags, hits=ph, rgb=pg)
            assert np.array_equal(bits(ph.numpy()), want_h) and np.array_equal(bits(pg.numpy()), want_g)
            ch, cg = gb.render_camera_ao(d_tris, cam, w, h, spp, seed, radius, flags, device="cuda")
            assert cg.is_cuda and np.array_equal(bits(ch.cpu().numpy()), want_h) and np.array_equal(bits(cg.cpu().numpy()), want_g)
            with pytest.raises(rt.RtError):
                gb.render_camera_ao(d_tris, cam, 0, h, spp, seed, radius)
            with pytest.raises(rt.RtError):
                gb.render_ao(d_tris.cpu().numpy(), pinned, spp, seed, radius)                   # tris must be on the device
            with pytest.raises(rt.RtError):
                gb.render_ao(d_tris, pinned, 65, seed, radius)
            with pytest.raises(rt.RtError):
                gb.render_ao(d_tris, pinned, spp, seed, radius, rt.RT_TRACE_ANY)


# ---- shading and the image ------------------------------------------------------------------------------
def test_shade_and_frame_match_the_oracle(rt, oracle, tmp_path):
    """§8(f).1: rt_shade bit-equal to the oracle's shade on the Cornell frame of the golden
    fixture (primary -> AO -> any-hit -> shade), and the PPM written from it has the pinned
    sha256 (tests/golden/cornell_96_frame.json)."""
    want = json.load(open(os.path.join(ROOT, "tests", "golden", "cornell_96_frame.json")))
    w, h, spp, seed = want["width"], want["height"], want["spp"], want["seed"]
    cam = rt.camera.cornell_camera(w, h)
    d_tris = rt.gen_scene(rt.RT_SCENE_CORNELL)
    tris = d_tris.cpu().numpy()
    with rt.Bvh.build(d_tris) as gb:
        rays = rt.gen_primary_rays(cam, w, h)
        hits = gb.trace(rays)
        radius = np.float32(0.05) * np.float32(np.linalg.norm(gb.bounds[3:] - gb.bounds[:3]))
        assert float(radius) == want["radius"]
        ao = rt.gen_ao_rays(d_tris, rays, hits, spp, seed, float(radius))
        occ = gb.trace(ao, rt.RT_TRACE_ANY)
        rgb = rt.shade(d_tris, rays, hits, occ, spp).cpu().numpy()
        rgb_plain = rt.shade(d_tris, rays, hits).cpu().numpy()
    o_rgb = oracle.shade(tris, rays.cpu().numpy(), hits.cpu().numpy(), occ.cpu().numpy(), spp)
    assert np.array_equal(bits(rgb), bits(o_rgb))
    assert np.array_equal(bits(rgb_plain), bits(oracle.shade(tris, rays.cpu().numpy(), hits.cpu().numpy())))
    assert hashlib.sha256(rgb.tobytes()).hexdigest() == want["rgb_sha256"]
    path = tmp_path / "frame.ppm"
    rt.image.write_ppm(str(path), rgb, w, h)
    data = path.read_bytes()
    assert len(data) == want["ppm_bytes"] and hashlib.sha256(data).hexdigest() == want["ppm_sha256"]
    # host inputs are staged by the library
    host = rt.shade(d_tris, rays.cpu().numpy(), hits.cpu().numpy(), occ.cpu().numpy(), spp)
    assert np.array_equal(bits(host), bits(o_rgb))


# ---- zero direction components ----------------------------------------------------------------------------
@pytest.mark.parametrize("wide", [False, True])
def test_zero_direction_rays_match_brute_force(rt, oracle, wide):
    """a ray with a +-0 direction component whose origin lies ON a box face was culled
    (1/d clamped to FLT_MAX gives 0 * FLT_MAX = 0 for the on-face plane).  Grid-aligned planes,
    origins on grid lines and vertices, every sign of zero: BVH2, BVH4 and the oracle against
    the all-pairs intersector, bit for bit."""
    tris, rays = zero_direction_cases()
    for key_bits in (30, 63):
        with rt.Bvh.build(dev(tris), key_bits=key_bits, wide=wide) as gb:
            for flags in (rt.RT_TRACE_WATERTIGHT, 0, rt.RT_TRACE_ANY | rt.RT_TRACE_WATERTIGHT):
                got = gb.trace(dev(rays), flags).cpu().numpy()
                want = oracle.brute(tris, rays, flags)
                assert np.array_equal(bits(got), bits(want)), (key_bits, flags)
                if not wide:
                    got = gb.trace(dev(np.tile(rays, (200, 1))), flags | rt.RT_TRACE_REORDER).cpu().numpy()
                    assert np.array_equal(bits(got), bits(np.tile(want, (200, 1)))), (key_bits, flags)
    # the advisor's two-triangle scene
    t = np.zeros((2, 12), dtype=np.float32)
    t[0] = [0, 0, 0, 0, 0.5, 0, 0, 0, 0.5, 0.5, 0, 0]
    t[1] = [2, 2, 2, 0, 3, 2, 2, 0, 3, 3, 2, 0]
    r = np.array([[0.5, 0.25, -1, 0, 0.0, 0, 1, np.inf], [0.5, 0.25, -1, 0, -0.0, 0, 1, np.inf]], dtype=np.float32)
    with rt.Bvh.build(t, wide=wide) as gb:
        got = gb.trace(r)
    assert oracle.hit_prims(got).tolist() == [0, 0] and got[:, 0].tolist() == [1.0, 1.0]


# ---- BVH files ------------------------------------------------------------------------------------------
def test_corrupt_bvh_files_are_rejected(rt, oracle, tmp_path):
    """bvh_load validates what it read.  Truncated files, out-of-range child ids, broken
    leaf ranges, primitive indices beyond n, a bad root and a tree deeper than the traversal
    stack all fail with RT_ERR_INVALID; the device is fine afterwards."""
    from gpu_raytracer_b200 import _lib
    tris = oracle.gen_scene(oracle.SCENE_SOUP, 5000, 0, seed=11)
    ob = oracle.OracleBvh(tris)
    rays = oracle.gen_random_rays(ob.bounds, 4, 0, 20000)
    want = ob.trace(rays)
    good = tmp_path / "good.bvh"
    with rt.Bvh.build(dev(tris)) as gb:
        gb.save(str(good))
    with rt.Bvh.load(str(good)) as lb:
        assert np.array_equal(bits(lb.trace(rays)), bits(want))
    blob = bytearray(good.read_bytes())
    n = 5000
    off_hdr, off_nodes = 24, 24 + 64
    off_tris = off_nodes + (n - 1) * 64

    def load_patched(mutate, name):
        b = bytearray(blob)
        b = mutate(b) or b
        p = tmp_path / (name + ".bvh")
        p.write_bytes(bytes(b))
        with pytest.raises(rt.RtError) as e:
            rt.Bvh.load(str(p))
        assert e.value.code == _lib.RT_ERR_INVALID, name

    def put_u32(b, off, v):
        b[off:off + 4] = int(v & 0xFFFFFFFF).to_bytes(4, "little")

    load_patched(lambda b: b[:len(b) // 2], "truncated")
    load_patched(lambda b: b + b"\0" * 16, "too_long")
    load_patched(lambda b: put_u32(b, off_nodes + 100 * 64 + 12, 0x7FFFFF00), "child_out_of_range")
    load_patched(lambda b: put_u32(b, off_nodes + 100 * 64 + 12, ~(n + 5)), "leaf_out_of_range")
    load_patched(lambda b: put_u32(b, off_nodes + 200 * 64 + 28, 4999), "range_L_beyond_split")
    load_patched(lambda b: put_u32(b, off_nodes + 200 * 64 + 60, n + 7), "range_R_beyond_n")
    load_patched(lambda b: put_u32(b, off_tris + 333 * 48 + 12, n), "prim_beyond_n")
    load_patched(lambda b: put_u32(b, off_hdr + 36, n + 1), "root_out_of_range")
    load_patched(lambda b: put_u32(b, off_hdr + 32, 31), "key_bits")
    load_patched(lambda b: put_u32(b, off_hdr + 12, n - 1), "header_n")
    load_patched(lambda b: put_u32(b, 16, 17), "meta_key_bits")

    # a valid but degenerate tree, 200 levels deep (a chain): every node's right child is a leaf
    m = 201
    hdr = np.zeros(16, dtype=np.uint32)
    hdr[3], hdr[8], hdr[9] = m, 30, m - 2          # n, key bits, root = last split
    nodes = np.zeros((m - 1, 16), dtype=np.uint32)
    for g in range(m - 1):
        nodes[g, 3] = (g - 1) if g > 0 else (~0 & 0xFFFFFFFF)   # child 0: node g-1, or leaf 0
        nodes[g, 7] = 0                                         # L
        nodes[g, 11] = (~(g + 1)) & 0xFFFFFFFF                  # child 1: leaf g+1
        nodes[g, 15] = g + 1                                    # R
    st = np.zeros((m, 12), dtype=np.uint32)
    st[:, 3] = np.arange(m)
    chain = tmp_path / "chain.bvh"
    with open(chain, "wb") as f:
        f.write(b"RTBVH1\0\0" + np.uint64(m).tobytes() + np.array([30, 0], dtype=np.uint32).tobytes())
        f.write(hdr.tobytes() + nodes.tobytes() + st.tobytes())
    with pytest.raises(rt.RtError) as e:
        rt.Bvh.load(str(chain))
    assert e.value.code == _lib.RT_ERR_INVALID and "deep" in str(e.value)
    # the same chain 60 levels deep is accepted: the check is the depth, not the shape
    m2 = 61
    hdr[3], hdr[9] = m2, m2 - 2
    with open(chain, "wb") as f:
        f.write(b"RTBVH1\0\0" + np.uint64(m2).tobytes() + np.array([30, 0], dtype=np.uint32).tobytes())
        f.write(hdr.tobytes() + nodes[:m2 - 1].tobytes() + st[:m2].tobytes())
    with rt.Bvh.load(str(chain)) as lb:
        assert lb.n_tris == m2
        assert (oracle.hit_prims(lb.trace(rays[:1000])) == -1).all()   # all-zero boxes: nothing to hit
    # and the device still works
    with rt.Bvh.load(str(good)) as lb:
        assert np.array_equal(bits(lb.trace(rays)), bits(want))
    ob.close()


# ---- runtime --------------------------------------------------------------------------------------------
def test_trim_and_pool_knob(rt, oracle):
    tris = dev(oracle.gen_scene(oracle.SCENE_SOUP, 200000, 0, seed=2))
    rt.Bvh.build(tris).close()
    rt.trim()
    rt.set_param("pool.keep_mb", 64)
    try:
        with rt.Bvh.build(tris) as gb:
            n = gb.n_tris
        assert n == 200000
    finally:
        rt.set_param("pool.keep_mb", 0)
    rt.trim(0)
    with pytest.raises(rt.RtError):
        rt.trim(1000)


def test_knobs_may_change_while_other_threads_trace(rt, oracle):
    """rt_set_param is documented as safe against concurrent calls: the knobs that only choose
    a launch shape are flipped while four threads trace; every result has the oracle's bits."""
    tris = oracle.gen_scene(oracle.SCENE_SOUP, 20000, 0, seed=1)
    ob = oracle.OracleBvh(tris)
    rays = [oracle.gen_random_rays(ob.bounds, 4, k * 40000, 40000) for k in range(4)]
    want = [ob.trace(r) for r in rays]
    stop = threading.Event()
    bad = []

    def flip():
        i = 0
        while not stop.is_set():
            rt.set_param("trace.chunk", (0, 32, 64, 128)[i % 4])
            rt.set_param("trace.leaf_min", (0, 4, 8, 16)[i % 4])
            rt.set_param("trace.variant", (1, 0)[i % 2])
            i += 1

    with rt.Bvh.build(dev(tris)) as gb:
        def work(k):
            d = dev(rays[k])
            for _ in range(20):
                got = gb.trace(d).cpu().numpy()
                if not np.array_equal(bits(got), bits(want[k])):
                    bad.append(k)
        f = threading.Thread(target=flip)
        f.start()
        ts = [threading.Thread(target=work, args=(k,)) for k in range(4)]
        [t.start() for t in ts]
        [t.join() for t in ts]
        stop.set()
        f.join()
    for name in ("trace.chunk", "trace.leaf_min"):
        rt.set_param(name, 0)
    rt.set_param("trace.variant", 1)
    assert not bad
    ob.close()


def test_frames_rendered_from_several_host_threads_at_once(rt):
    """rt_render_ao keeps its streams, slots and counters per host thread: four threads render
    different frames against one handle at the same time (pinned host buffers, compact nodes in
    two of them), each gets the bytes of its own sequential frame."""
    d_tris = rt.gen_scene(rt.RT_SCENE_SPHEREFLAKE, 3, 4, 0)
    w, h, spp = 1280, 900, 2                                   # 1.15M rays: several bands
    with rt.Bvh.build(d_tris, compact=True) as gb:
        radius = float(np.float32(0.05) * np.float32(np.linalg.norm(gb.bounds[3:] - gb.bounds[:3])))
        cams = [rt.camera.frame_box(gb.bounds, w, h, direction=d) for d in ((0.3, 0.25, -1.0), (1.0, 0.2, 0.1), (-0.5, 0.8, 0.4), (0.0, -1.0, 0.2))]
        rays = [rt.gen_primary_rays(c, w, h).cpu().pin_memory() for c in cams]
        want = [gb.render_ao(d_tris, r, spp, 11 + k, radius) for k, r in enumerate(rays)]
        want = [(a.clone(), b.clone()) for a, b in want]
        got = [None] * 4

        def work(k):
            for _ in range(5):
                got[k] = gb.render_ao(d_tris, rays[k], spp, 11 + k, radius)
        ts = [threading.Thread(target=work, args=(k,)) for k in range(4)]
        [t.start() for t in ts]
        [t.join() for t in ts]
        for k in range(4):
            assert torch.equal(got[k][0].view(torch.int32), want[k][0].view(torch.int32)), k
            assert torch.equal(got[k][1].view(torch.int32), want[k][1].view(torch.int32)), k
        assert (rt.hit_prims(want[0][0]) >= 0).float().mean() > 0.05


def test_c99_program_runs_on_the_device(rt, tmp_path):
    exe = tmp_path / "smoke"
    lib_dir = os.path.dirname(rt.SO_PATH)
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Wextra", "-Werror", "-pedantic", "-I", os.path.join(ROOT, "include"),
                    os.path.join(ROOT, "tests", "c99", "smoke.c"), "-o", str(exe), rt.SO_PATH,
                    f"-Wl,-rpath,{lib_dir}", "-lm"], check=True)
    res = subprocess.run([str(exe)], capture_output=True, text=True)
    assert res.returncode == 0 and res.stdout.strip() == "ok", res.stdout + res.stderr


def test_cpp_bench_main_gives_the_python_paths_hits(rt, tmp_path):
    """tools/cpp/bench_main.cpp (SURVEY.md §8b "a small C++ bench main"): no Python, no PyTorch,
    cudaMalloc'd buffers -> rt_gen_scene / build_bvh / trace_rays.  Same camera, same hit bytes."""
    exe = tmp_path / "bench_main"
    lib_dir = os.path.dirname(rt.SO_PATH)
    cuda = os.environ.get("CUDA_HOME", "/usr/local/cuda")
    subprocess.run(["g++", "-std=c++17", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), "-I", os.path.join(cuda, "include"),
                    os.path.join(ROOT, "tools", "cpp", "bench_main.cpp"), "-o", str(exe), rt.SO_PATH,
                    "-L", os.path.join(cuda, "lib64"), "-lcudart", f"-Wl,-rpath,{lib_dir}", f"-Wl,-rpath,{os.path.join(cuda, 'lib64')}"],
                   check=True)
    w, h = 320, 200
    cam = rt.camera.cornell_camera(w, h)
    res = subprocess.run([str(exe), "cornell", str(w), str(h), "2", "cam=" + ",".join(repr(float(v)) for v in cam)],
                         capture_output=True, text=True)
    assert res.returncode == 0, res.stdout + res.stderr
    out = json.loads(res.stdout)
    with rt.Bvh.build(rt.gen_scene(rt.RT_SCENE_CORNELL)) as gb:
        hits = gb.trace(rt.gen_primary_rays(cam, w, h)).cpu().numpy()
    fnv = 1469598103934665603
    for b in hits.tobytes():
        fnv = ((fnv ^ b) * 1099511628211) & 0xFFFFFFFFFFFFFFFF
    assert out["n_tris"] == 1024 and out["n_rays"] == w * h and out["host_path_equal"] is True
    assert out["hits"] == int((hits.view(np.int32)[:, 3] >= 0).sum()) > 0
    assert out["hits_fnv1a"] == f"{fnv:016x}"
    assert out["build_ms"] > 0 and out["trace_ms"] > 0 and out["kernel_launches"] > 10


# ---- sharded BUILD (§8f.4) -------------------------------------------------------------------------------
@pytest.mark.parametrize("parts", [1, 2, 3, 5])
def test_sharded_build_merges_to_the_single_bvh(rt, oracle, parts):
    """build_bvh_sharded / trace_rays_merged / rt_merge_hits: the SCENE cut into contiguous ranges
    of the triangle array, one BVH per range (all on GPU 0 here; the last on another GPU when
    there is one), every part traces all rays, hits merged by (t, prim + base): the single
    BVH's hits, bit for bit, for closest / any-hit / watertight."""
    devices = [0] * parts
    if torch.cuda.device_count() > 1 and parts > 1:
        devices[-1] = torch.cuda.device_count() - 1
    tris = oracle.gen_scene(oracle.SCENE_SPHEREFLAKE, 3, 4, 0)
    ob = oracle.OracleBvh(tris)
    rays = oracle.gen_random_rays(ob.bounds, 4, 0, 300_001)
    cam = rt.camera.frame_box(ob.bounds, 320, 200)
    rays = np.concatenate([rays, oracle.gen_primary_rays(cam, 320, 200)])
    with rt.ShardedBvh.build(tris, devices) as sb:
        from gpu_raytracer_b200 import sharding
        assert sb.prim_base == [sharding.tri_range(ob.n, r, parts)[0] for r in range(parts)]
        assert sum(h.n_tris for h in sb.handles) == ob.n
        for flags in (0, rt.RT_TRACE_ANY, rt.RT_TRACE_WATERTIGHT):
            want = ob.trace(rays, flags)
            # fused: every device merges its slice of the batch out of the parts' (peer) memory;
            # 0: the parts' hits are copied to the first device and merged there
            for fused in (1, 0):
                rt.set_param("merge.fused", fused)
                try:
                    got = sb.trace(rays, flags)
                finally:
                    rt.set_param("merge.fused", 1)
                assert np.array_equal(bits(got), bits(want)), (parts, flags, fused)
        # the exchange step alone: device hits of every part, merged by the kernel == numpy statement
        d_rays = dev(rays)
        local = [h.trace(d_rays) for h in sb.handles if h.info().device == 0]
        if len(local) == parts:
            merged = rt.merge_hits(local, sb.prim_base).cpu().numpy()
            assert np.array_equal(bits(merged), bits(ob.trace(rays)))
            host = sharding.merge_hits_host([x.cpu().numpy() for x in local], sb.prim_base)
            assert np.array_equal(bits(host), bits(merged))
    ob.close()
    with pytest.raises(rt.RtError):
        rt.ShardedBvh.build(dev(tris), [0])          # device triangles: each part uploads its own range


@pytest.mark.parametrize("shape", [256 * 16 + 2, 128 * 16 + 4, 64 * 16 + 4])
def test_walk_kernel_shapes_build_the_same_tree(rt, oracle, shape):
    """`build.walk` (threads * 16 + leaves per thread): the bottom-up pass with several consecutive
    leaves per thread (measured slower than one, profiles/r02k_walk_shapes.jsonl; kept as a knob)
    builds the oracle's tree bit for bit."""
    rt.set_param("build.walk", shape)
    try:
        for n, kb in [(2, 30), (3, 30), (255, 30), (257, 63), (1025, 30), (40001, 30), (200000, 63)]:
            t = oracle.gen_scene(oracle.SCENE_SOUP, n, 0, seed=n)
            ob = oracle.OracleBvh(t, kb)
            with rt.Bvh.build(dev(t), key_bits=kb) as b:
                l, r, p = b.read_topology()
                assert np.array_equal(l, ob.left) and np.array_equal(r, ob.right) and np.array_equal(p, ob.parent), n
                assert np.array_equal(bits(b.read_boxes()), bits(ob.boxes)), n
    finally:
        rt.set_param("build.walk", 0)
