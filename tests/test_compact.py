"""The 32-byte binary layout (csrc/compact.cu, RT_BUILD_COMPACT): 16-bit child boxes on one grid
per scene.

CPU: the oracle's grid and packing (oracle.compact_grid / compact_pack) keep one whole step
around every box, whatever the scene's place and size.  GPU: the device's nodes equal that
packing of its own 64-byte nodes bit for bit, and tracing through them returns the oracle's
(binary-tree) hits bit for bit: closest, any-hit, watertight (which stays on the 64-byte nodes),
reordered, zero direction components, far-away and degenerate scenes, after a refit, after a load.
"""
import numpy as np
import pytest

from conftest import bits
from test_host_logic import zero_direction_cases

SCENES = {"cornell": (0, 0, 0, 0), "soup20k": (1, 20000, 0, 1), "flake": (2, 2, 4, 0), "terrain": (3, 64, 16, 3)}


def planes(words, grid):
    """float64 [m, 6] boxes that the packed words stand for"""
    lo = (words & 0xFFFF).astype(np.float64)
    hi = (words >> 16).astype(np.float64)
    o, s = grid[:3].astype(np.float64), grid[3:].astype(np.float64)
    return np.concatenate([o + lo * s, o + hi * s], axis=1)


@pytest.mark.parametrize("shift,scale", [(0.0, 1.0), (1000.0, 1.0), (-3.0e6, 2.0), (0.25, 1.0e-4), (5.0, 1.0e5)])
def test_packing_keeps_a_whole_step_around_every_box(oracle, shift, scale):
    rng = np.random.default_rng(5)
    lo = rng.random((5000, 3))
    boxes = np.concatenate([lo, lo + rng.random((5000, 3)) * 0.01], axis=1) * scale + shift
    boxes = boxes.astype(np.float32)
    boxes[0, 3:] = boxes[0, :3]                                   # a point
    bounds = np.concatenate([boxes[:, :3].min(0), boxes[:, 3:].max(0)])
    grid = oracle.compact_grid(bounds)
    assert (grid[3:] > 0).all()
    words = oracle.compact_pack(boxes, grid)
    got = planes(words, grid)
    step = grid[3:].astype(np.float64)
    b = boxes.astype(np.float64)
    assert (got[:, :3] <= b[:, :3] - step * (1 - 1e-9)).all() and (got[:, 3:] >= b[:, 3:] + step * (1 - 1e-9)).all()
    assert (got[:, :3] >= b[:, :3] - 2.01 * step).all() and (got[:, 3:] <= b[:, 3:] + 2.01 * step).all(), "and no more than two"
    assert ((words & 0xFFFF) >= 1).all() and ((words >> 16) <= 65534).all(), "never clamped"


def test_grid_of_a_flat_scene(oracle):
    grid = oracle.compact_grid(np.array([0, 0, 2, 1, 1, 2], dtype=np.float32))   # no extent on z
    assert grid[5] > 0
    words = oracle.compact_pack(np.array([[0, 0, 2, 1, 1, 2]], dtype=np.float32), grid)
    got = planes(words, grid)[0]
    assert got[2] < 2 < got[5]


# ---- GPU -------------------------------------------------------------------------------------
gpu = pytest.mark.gpu


def dev(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


def raw_nodes(gb):
    """the handle's 64-byte nodes as float32 [n - 1, 16] (and the same bytes as int32)"""
    from gpu_raytracer_b200.sharding import _as_tensor
    ptr, nbytes = gb.device_buffers()[0]
    raw = _as_tensor(ptr, nbytes).cpu().numpy()
    return raw.view(np.float32).reshape(-1, 16), raw.view(np.int32).reshape(-1, 16)


@gpu
@pytest.mark.parametrize("key_bits", [30, 63])
@pytest.mark.parametrize("name", list(SCENES))
def test_device_nodes_equal_the_oracles_packing(rt, oracle, name, key_bits):
    kind, p0, p1, seed = SCENES[name]
    tris = oracle.gen_scene(kind, p0, p1, seed)
    with rt.Bvh.build(dev(tris), key_bits=key_bits, compact=True) as gb:
        nodes, grid = gb.read_compact()
        f, i = raw_nodes(gb)
        n = gb.n_tris
        bounds = gb.bounds
    f, i = f[:n - 1], i[:n - 1]
    assert np.array_equal(bits(grid), bits(oracle.compact_grid(bounds)))
    for c in (0, 1):
        box = np.concatenate([f[:, 8 * c:8 * c + 3], f[:, 8 * c + 4:8 * c + 7]], axis=1)
        assert np.array_equal(nodes["box"][:, c, :], oracle.compact_pack(box, grid)), c
        assert np.array_equal(nodes["child"][:, c], i[:, 8 * c + 3]), "child ids are the 64-byte node's"


@gpu
@pytest.mark.parametrize("name", list(SCENES))
def test_compact_hits_are_bit_identical(rt, oracle, name):
    kind, p0, p1, seed = SCENES[name]
    tris = oracle.gen_scene(kind, p0, p1, seed)
    v = tris.reshape(-1, 4)[:, :3]
    box = np.concatenate([v.min(0), v.max(0)])
    rays = oracle.gen_random_rays(box, 4, 0, 150001)
    cam = rt.camera.frame_box(box, 160, 120)
    rays = np.concatenate([rays, oracle.gen_primary_rays(cam, 160, 120)])
    rays[7, 4:7] = 0.0
    rays[11, 0:3] = box[3:] * 4
    rays[12, 0:3] = box[3:] * 1.0e4                                # an origin 10^4 scene sizes away
    rays[13, 0:3] = box[:3]                                        # on the corner of the scene box
    rays[14:20, 4] = 0.0                                           # parallel to a grid axis
    d_rays = dev(rays)
    for key_bits in (30, 63):
        ob = oracle.OracleBvh(tris, key_bits)
        with rt.Bvh.build(dev(tris), key_bits=key_bits, compact=True) as gb:
            for flags in (0, rt.RT_TRACE_ANY, rt.RT_TRACE_WATERTIGHT, rt.RT_TRACE_ANY | rt.RT_TRACE_WATERTIGHT):
                want = ob.trace(rays, flags)
                got = gb.trace(d_rays, flags).cpu().numpy()
                assert np.array_equal(bits(got), bits(want)), (key_bits, flags)
                got = gb.trace(d_rays, flags | rt.RT_TRACE_REORDER).cpu().numpy()
                assert np.array_equal(bits(got), bits(want)), (key_bits, flags, "reorder")
            got = gb.trace(rays)                                   # host buffers
            assert np.array_equal(bits(np.asarray(got)), bits(ob.trace(rays)))
            hits, n_nodes, n_tests = gb.trace_stats(d_rays)
            assert np.array_equal(bits(hits.cpu().numpy()), bits(ob.trace(rays)))
            _, (o_nodes, o_tests) = ob.trace(rays, counters=True)
            # boxes one to two steps larger: a few more visits than the exact boxes, never fewer
            assert o_nodes <= n_nodes < 1.1 * o_nodes and o_tests <= n_tests < 1.5 * o_tests
            rt.set_param("trace.compact", 0)
            try:
                _, n_nodes, n_tests = gb.trace_stats(d_rays)
            finally:
                rt.set_param("trace.compact", 1)
            assert (n_nodes, n_tests) == (o_nodes, o_tests), "the knob switches back to the 64-byte nodes"
        ob.close()


@gpu
def test_compact_zero_direction_rays_match_brute_force(rt, oracle):
    tris, rays = zero_direction_cases()
    with rt.Bvh.build(dev(tris), compact=True) as gb:
        for flags in (0, rt.RT_TRACE_ANY, rt.RT_TRACE_WATERTIGHT):
            got = gb.trace(dev(rays), flags).cpu().numpy()
            assert np.array_equal(bits(got), bits(oracle.brute(tris, rays, flags))), flags


@gpu
@pytest.mark.parametrize("shift,scale", [(1000.0, 1.0), (-3.0e6, 2.0), (0.25, 1.0e-4), (5.0, 1.0e5)])
def test_compact_far_and_tiny_scenes(rt, oracle, shift, scale):
    """scenes whose coordinates are large against their size (the grid is as coarse as floats
    there), tiny, and huge: same hits as the oracle's binary tree"""
    tris = oracle.gen_scene(oracle.SCENE_SOUP, 3000, 0, seed=7)
    tris[:, [0, 1, 2, 4, 5, 6, 8, 9, 10]] = (tris[:, [0, 1, 2, 4, 5, 6, 8, 9, 10]] * scale + shift).astype(np.float32)
    v = tris.reshape(-1, 4)[:, :3]
    box = np.concatenate([v.min(0), v.max(0)])
    rays = oracle.gen_random_rays(box, 9, 0, 40000)
    want = oracle.OracleBvh(tris).trace(rays)
    assert (oracle.hit_prims(want) >= 0).mean() > 0.005
    with rt.Bvh.build(dev(tris), compact=True) as gb:
        assert np.array_equal(bits(gb.trace(dev(rays)).cpu().numpy()), bits(want))
        assert np.array_equal(bits(gb.trace(dev(rays), rt.RT_TRACE_ANY).cpu().numpy()),
                              bits(oracle.OracleBvh(tris).trace(rays, oracle.TRACE_ANY)))


@gpu
def test_compact_edge_cases_refit_and_load(rt, oracle, tmp_path):
    base = oracle.gen_scene(oracle.SCENE_SOUP, 5000, 0, seed=3)
    ray = oracle.gen_random_rays(np.array([0, 0, 0, 1, 1, 1], dtype=np.float32), 1, 0, 20000)
    for n in (0, 1):                                                # too small: the binary tree answers
        with rt.Bvh.build(base[:n], compact=True) as gb:
            got = gb.trace(ray)
            if n:
                assert np.array_equal(bits(got), bits(oracle.OracleBvh(base[:n]).trace(ray)))
            else:
                assert (oracle.hit_prims(got) == -1).all()
            with pytest.raises(rt.RtError):
                gb.read_compact()
            with pytest.raises(rt.RtError):
                gb.compact()
    for n in (2, 3, 9, 33):
        with rt.Bvh.build(base[:n], compact=True) as gb:
            assert gb.read_compact()[0].shape[0] == n - 1
            assert np.array_equal(bits(gb.trace(ray)), bits(oracle.OracleBvh(base[:n]).trace(ray)))
    dup = np.repeat(base[:3], 50, axis=0)                           # equal centroids: the deepest trees
    with rt.Bvh.build(dup, key_bits=63, compact=True) as gb:
        assert np.array_equal(bits(gb.trace(ray)), bits(oracle.OracleBvh(dup, 63).trace(ray)))
    flat = base.copy()
    flat[:, [2, 6, 10]] = np.float32(0.5)                           # a scene with no extent on z
    with rt.Bvh.build(flat, compact=True) as gb:
        assert np.array_equal(bits(gb.trace(ray)), bits(oracle.OracleBvh(flat).trace(ray)))
    # refit: moved vertices, same topology, the nodes are packed again on the new scene box's grid
    moved = base.copy()
    moved[:, [1, 5, 9]] += np.float32(0.25) * np.sin(moved[:, [0, 4, 8]] * 7).astype(np.float32)
    with rt.Bvh.build(dev(base), compact=True) as gb:
        before = gb.read_compact()[1]
        gb.refit(dev(moved))
        nodes, grid = gb.read_compact()
        assert not np.array_equal(before, grid)
        assert np.array_equal(bits(grid), bits(oracle.compact_grid(gb.bounds)))
        rt.set_param("trace.compact", 0)
        try:
            want = gb.trace(dev(ray)).cpu().numpy()
        finally:
            rt.set_param("trace.compact", 1)
        assert np.array_equal(bits(gb.trace(dev(ray)).cpu().numpy()), bits(want))
        assert np.array_equal(bits(want), bits(oracle.brute(moved, ray)))
        # a saved BVH holds the 64-byte nodes; the loaded one is packed on request
        path = str(tmp_path / "soup.bvh")
        gb.save(path)
    with rt.Bvh.load(path) as lb:
        with pytest.raises(rt.RtError):
            lb.read_compact()
        lb.compact()
        assert np.array_equal(lb.read_compact()[0], nodes)
        assert np.array_equal(bits(lb.trace(dev(ray)).cpu().numpy()), bits(want))
