"""GPU parity tests (SURVEY.md §4 rungs T2-T7): the CUDA path, called through the
C ABI (ctypes), against the CPU oracle on the same inputs.

Bar (BASELINE.json:5): bit-exact Morton codes, sorted order, topology, primitive
indices and node boxes; identical hit/miss and primitive ids; t/u/v within 1e-5
relative.  Because both sides use the same IEEE operations in the same order,
hits are in fact asserted bit-exact here; the 1e-5 check is kept as the formal
tolerance.
"""
import threading

import numpy as np
import pytest

from conftest import bits
from helpers import check_boxes, check_tree_invariants, compare_hits

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


@pytest.fixture(scope="module", autouse=True)
def _need_cuda(rt):
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    assert rt.lib().rt_device_count() >= 1


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


SCENES = {
    "cornell": (0, 0, 0, 0),
    "soup20k": (1, 20000, 0, 1),
    "flake": (2, 2, 4, 0),
    "terrain": (3, 64, 16, 3),
}


# ---- generators: device bytes == oracle bytes ------------------------------------
@pytest.mark.parametrize("name", list(SCENES) + ["flake_deep", "terrain_big"])
def test_scene_generators_match_oracle(rt, oracle, name):
    kind, p0, p1, seed = {**SCENES, "flake_deep": (2, 4, 2, 0), "terrain_big": (3, 300, 0, 5)}[name]
    want = oracle.gen_scene(kind, p0, p1, seed)
    got = rt.gen_scene(kind, p0, p1, seed).cpu().numpy()
    assert got.shape == want.shape
    assert np.array_equal(bits(got), bits(want))
    # sub-range into a host buffer
    part = rt.gen_scene(kind, p0, p1, seed, first=100, count=333, device=None)
    assert np.array_equal(bits(part), bits(want[100:433]))


def test_ray_generators_match_oracle(rt, oracle):
    cam = rt.camera.cornell_camera(256, 256)
    want = oracle.gen_primary_rays(cam, 256, 256)
    got = rt.gen_primary_rays(cam, 256, 256).cpu().numpy()
    assert np.array_equal(bits(got), bits(want))
    box = np.array([-1, 0, 2, 3, 0.5, 7], dtype=np.float32)
    want = oracle.gen_random_rays(box, 4, 12345, 50000)
    got = rt.gen_random_rays(box, 4, 12345, 50000).cpu().numpy()
    assert np.array_equal(bits(got), bits(want))


# ---- T2-T4: build ------------------------------------------------------------------
def _build_and_compare(rt, oracle, tris, key_bits, timings=True):
    ob = oracle.OracleBvh(tris, key_bits)
    with rt.Bvh.build(dev(tris), key_bits=key_bits, keep_keys=True, timings=timings) as gb:
        n = gb.n_tris
        assert n == ob.n
        assert np.array_equal(bits(gb.bounds), bits(ob.bounds)), "A1 scene bounds"
        keys, idx = gb.read_sorted()
        assert np.array_equal(keys, ob.keys), "A2/A3 sorted Morton codes"
        assert np.array_equal(idx, ob.idx), "A3 sorted primitive indices"
        left, right, parent = gb.read_topology()
        assert np.array_equal(left, ob.left) and np.array_equal(right, ob.right), "A4 children"
        assert np.array_equal(parent, ob.parent), "A4 parents"
        boxes = gb.read_boxes()
        assert np.array_equal(bits(boxes), bits(ob.boxes)), "A5 node boxes"
        check_tree_invariants(n, left, right, parent, idx)
        check_boxes(n, left, right, boxes, tris, idx)
        if timings:
            t = gb.build_times()
            assert t["total"] > 0
    return ob


@pytest.mark.parametrize("key_bits", [30, 63])
@pytest.mark.parametrize("name", list(SCENES))
def test_build_matches_oracle(rt, oracle, name, key_bits):
    kind, p0, p1, seed = SCENES[name]
    _build_and_compare(rt, oracle, oracle.gen_scene(kind, p0, p1, seed), key_bits)


@pytest.mark.parametrize("key_bits", [30, 63])
def test_build_ragged_sizes(rt, oracle, key_bits):
    """Sizes around the sort tile and warp boundaries, host-pointer input."""
    for n in [2, 3, 31, 32, 33, 255, 256, 257, 5119, 5120, 5121, 7679, 7680, 7681, 15361, 40001]:
        tris = oracle.gen_scene(oracle.SCENE_SOUP, n, 0, seed=n)
        ob = oracle.OracleBvh(tris, key_bits)
        with rt.Bvh.build(tris, key_bits=key_bits, keep_keys=True) as gb:  # numpy -> staged H2D
            keys, idx = gb.read_sorted()
            assert np.array_equal(keys, ob.keys) and np.array_equal(idx, ob.idx), n
            left, right, parent = gb.read_topology()
            assert np.array_equal(left, ob.left) and np.array_equal(right, ob.right), n
            assert np.array_equal(parent, ob.parent), n
            assert np.array_equal(bits(gb.read_boxes()), bits(ob.boxes)), n


def test_duplicate_codes(rt, oracle):
    """Many triangles per Morton cell: the index tie-break (A4) must hold."""
    base = oracle.gen_scene(oracle.SCENE_SOUP, 300, 0, seed=2)
    tris = np.repeat(base, 40, axis=0)  # 40 exact copies of each triangle
    ob = _build_and_compare(rt, oracle, tris, 30)
    assert len(np.unique(ob.keys)) <= 300
    _build_and_compare(rt, oracle, tris, 63)
    # all centroids identical
    one = np.repeat(base[:1], 1000, axis=0)
    _build_and_compare(rt, oracle, one, 30)


# ---- T5: hits -------------------------------------------------------------------------
def _trace_both(rt, oracle, tris, rays, flags=0, key_bits=30):
    ob = oracle.OracleBvh(tris, key_bits)
    want = ob.trace(rays, flags)
    with rt.Bvh.build(dev(tris), key_bits=key_bits) as gb:
        got = gb.trace(dev(rays), flags).cpu().numpy()
    return got, want


def test_config1_cornell_hits(rt, oracle):
    """Config 1 (BASELINE.json:7): Cornell box, 256x256 primary rays."""
    tris = oracle.gen_scene(oracle.SCENE_CORNELL)
    rays = oracle.gen_primary_rays(rt.camera.cornell_camera(256, 256), 256, 256)
    got, want = _trace_both(rt, oracle, tris, rays)
    compare_hits(got, want, rtol=1e-5)
    assert np.array_equal(bits(got), bits(want))
    assert (oracle.hit_prims(want) >= 0).sum() > 50000
    # and against brute force (T0), with the oracle's own tolerance for grazing rays
    hb = oracle.brute(tris, rays)
    assert np.array_equal(bits(got), bits(hb)), "BVH traversal == all-pairs intersector, exactly"


@pytest.mark.parametrize("key_bits", [30, 63])
@pytest.mark.parametrize("name", ["soup20k", "flake", "terrain"])
def test_random_ray_hits(rt, oracle, name, key_bits):
    kind, p0, p1, seed = SCENES[name]
    tris = oracle.gen_scene(kind, p0, p1, seed)
    v = tris.reshape(-1, 4)[:, :3]
    box = np.concatenate([v.min(0), v.max(0)])
    rays = oracle.gen_random_rays(box, 4, 0, 100000)
    got, want = _trace_both(rt, oracle, tris, rays, 0, key_bits)
    compare_hits(got, want, rtol=1e-5)
    assert np.array_equal(bits(got), bits(want))
    assert (oracle.hit_prims(want) >= 0).sum() > 1000
    got, want = _trace_both(rt, oracle, tris, rays, 1, key_bits)
    assert np.array_equal(bits(got), bits(want)), "any-hit"


def test_ao_pipeline(rt, oracle):
    """primary -> hits -> 4 AO rays per pixel -> any-hit; every stage compared."""
    tris = oracle.gen_scene(oracle.SCENE_SPHEREFLAKE, 2, 6)
    v = tris.reshape(-1, 4)[:, :3]
    box = np.concatenate([v.min(0), v.max(0)])
    cam = rt.camera.frame_box(box, 192, 128, fill=1.8)
    ob = oracle.OracleBvh(tris)
    rays = oracle.gen_primary_rays(cam, 192, 128)
    hits = ob.trace(rays)
    radius = 0.05 * float(np.linalg.norm(box[3:] - box[:3]))
    ao = oracle.gen_ao_rays(tris, rays, hits, 4, 2, radius)
    occ = ob.trace(ao, 1)
    d_tris = dev(tris)
    with rt.Bvh.build(d_tris) as gb:
        g_rays = rt.gen_primary_rays(cam, 192, 128)
        g_hits = gb.trace(g_rays)
        assert np.array_equal(bits(g_hits.cpu().numpy()), bits(hits))
        g_ao = rt.gen_ao_rays(d_tris, g_rays, g_hits, 4, 2, radius)
        assert np.array_equal(bits(g_ao.cpu().numpy()), bits(ao))
        g_occ = gb.trace(g_ao, 1)
        assert np.array_equal(bits(g_occ.cpu().numpy()), bits(occ))
        # the instrumented kernel returns the same hits and the oracle's counters
        s_hits, nn, nt = gb.trace_stats(g_rays)
        assert np.array_equal(bits(s_hits.cpu().numpy()), bits(hits))
        assert (nn, nt) == ob.trace(rays, counters=True)[1]
        rgb = rt.shade(d_tris, g_rays, g_hits, g_occ, 4).cpu().numpy()
        assert rgb.shape == (192 * 128, 3) and np.isfinite(rgb).all()
        assert rgb.min() >= 0 and rgb.max() <= 1.0 + 1e-6 and rgb.max() > 0.5
    assert (oracle.hit_prims(hits) >= 0).mean() > 0.3


def test_tie_break_and_ranges(rt, oracle):
    tri = np.array([[0, 0, 1, 0, 1, 0, 1, 0, 0, 1, 1, 0]], dtype=np.float32)
    far = np.array([[0, 0, 2, 0, 1, 0, 2, 0, 0, 1, 2, 0]], dtype=np.float32)
    tris = np.concatenate([far, tri, tri, tri])
    rays = np.array([[0.25, 0.25, 0, 0, 0, 0, 1, np.inf],
                     [0.25, 0.25, 0, 1.0, 0, 0, 1, np.inf],
                     [0.25, 0.25, 0, 0, 0, 0, 1, 1.0]], dtype=np.float32)
    got, want = _trace_both(rt, oracle, tris, rays)
    assert np.array_equal(bits(got), bits(want))
    assert oracle.hit_prims(got).tolist() == [1, 0, -1]


# ---- T7: edge cases ---------------------------------------------------------------------
def test_edge_cases(rt, oracle):
    ray = np.array([[0.25, 0.25, -1, 0, 0, 0, 1, np.inf]], dtype=np.float32)
    tri = np.array([[0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0]], dtype=np.float32)
    # N = 0: valid empty handle, every ray misses
    with rt.Bvh.build(np.zeros((0, 12), dtype=np.float32)) as gb:
        assert gb.n_tris == 0
        h = gb.trace(ray)
        assert oracle.hit_prims(h)[0] == -1 and np.isinf(h[0, 0])
    # N = 1 and N = 2
    for tris in (tri, np.concatenate([tri, tri + np.float32(3)])):
        got, want = _trace_both(rt, oracle, tris, ray)
        assert np.array_equal(bits(got), bits(want))
        _build_and_compare(rt, oracle, tris, 30)
        _build_and_compare(rt, oracle, tris, 63)
    # degenerate triangle + flat scene + zero direction components + origin on a face
    deg = tri.copy(); deg[0, 8:11] = deg[0, 4:7]
    got, want = _trace_both(rt, oracle, np.concatenate([deg, tri]), ray)
    assert np.array_equal(bits(got), bits(want)) and oracle.hit_prims(got)[0] == 1
    flat = oracle.gen_scene(oracle.SCENE_CORNELL)[:128]
    rays = np.array([[0.3, 1, 0.3, 0, 0, -1, 0, np.inf], [0.3, 0, 0.3, 0, 0, -1, 0, np.inf],
                     [0.0, 1, 0.3, 0, 0, -1, 0, np.inf], [0.3, 1, 0.3, 0, -0.0, -1, -0.0, np.inf]],
                    dtype=np.float32)
    got, want = _trace_both(rt, oracle, flat, rays)
    assert np.array_equal(bits(got), bits(want))
    _build_and_compare(rt, oracle, flat, 30)
    # zero rays
    with rt.Bvh.build(tri) as gb:
        assert gb.trace(np.zeros((0, 8), dtype=np.float32)).shape == (0, 4)


# ---- sort alone (A3) ----------------------------------------------------------------
@pytest.mark.parametrize("variant", [0, 1, 2, 3])
@pytest.mark.parametrize("bits", [0, 8, 9, 10, 11])
@pytest.mark.parametrize("key_bits", [30, 63])
def test_sort_pairs_is_stable(rt, variant, bits, key_bits):
    """Every digit width and tile shape of the onesweep kernel against numpy's stable
    argsort: uniform keys (optimistic ranking path), few distinct keys and one single
    key (ballot fallback; stability decides the order)."""
    rt.set_param("sort.variant", variant)
    rt.set_param("sort.bits", bits)
    try:
        rng = np.random.default_rng(variant + key_bits + 100 * bits)
        for n, distinct in [(1, 0), (1000, 0), (7680, 12), (7681, 3), (100003, 12), (1 << 20, 0),
                            (300001, 1), (1 << 20, 5)]:
            hi = 1 << (distinct if distinct else key_bits)
            if key_bits == 30:
                keys = rng.integers(0, min(hi, 1 << 30), size=n, dtype=np.uint32)
                if distinct:
                    keys = keys * np.uint32(0x01041041 >> 2)  # spread over all digits
                    keys &= np.uint32((1 << 30) - 1)
                tk = torch.from_numpy(keys.view(np.int32)).cuda()
            else:
                keys = rng.integers(0, min(hi, 1 << 63), size=n, dtype=np.uint64)
                if distinct:
                    keys = (keys * np.uint64(0x0040100401004010)) & np.uint64((1 << 63) - 1)
                tk = torch.from_numpy(keys.view(np.int64)).cuda()
            vals = np.arange(n, dtype=np.int32)
            tv = torch.from_numpy(vals).cuda()
            rt.sort_pairs(tk, tv, key_bits)
            order = np.argsort(keys, kind="stable")
            assert np.array_equal(tv.cpu().numpy(), order.astype(np.int32)), (n, distinct)
            assert np.array_equal(tk.cpu().numpy().view(keys.dtype), keys[order]), (n, distinct)
    finally:
        rt.set_param("sort.variant", 0)
        rt.set_param("sort.bits", 0)


# ---- full-size properties (config 2) ------------------------------------------------------
def test_config2_million_triangles(rt, oracle):
    """Config 2 (BASELINE.json:8): 1M-triangle soup, 1920x1080 primary rays."""
    n = 1000000
    d_tris = rt.gen_scene(rt.RT_SCENE_SOUP, n, 0, 1)
    tris = d_tris.cpu().numpy()
    assert np.array_equal(bits(tris[:5000]), bits(oracle.gen_scene(oracle.SCENE_SOUP, n, 0, 1, 0, 5000)))
    ob = oracle.OracleBvh(tris)
    with rt.Bvh.build(d_tris, keep_keys=True) as gb:
        keys, idx = gb.read_sorted()
        assert (np.diff(keys.astype(np.int64)) >= 0).all(), "sortedness"
        assert np.array_equal(np.sort(idx), np.arange(n, dtype=np.uint32)), "permutation"
        assert np.array_equal(keys, ob.keys) and np.array_equal(idx, ob.idx)
        left, right, parent = gb.read_topology()
        assert np.array_equal(left, ob.left) and np.array_equal(right, ob.right)
        assert np.array_equal(parent, ob.parent)
        assert np.array_equal(bits(gb.read_boxes()), bits(ob.boxes))
        cam = rt.camera.frame_box(gb.bounds, 1920, 1080)
        g_rays = rt.gen_primary_rays(cam, 1920, 1080)
        got = gb.trace(g_rays).cpu().numpy()
        want = ob.trace(g_rays.cpu().numpy())
        compare_hits(got, want)
        assert np.array_equal(bits(got), bits(want))
        assert (oracle.hit_prims(got) >= 0).mean() > 0.2
        # host-buffer path (chunked H2D / trace / D2H pipeline) gives the same bytes
        host_rays = g_rays.cpu().numpy()
        got2 = gb.trace(host_rays)
        assert isinstance(got2, np.ndarray) and np.array_equal(bits(got2), bits(want))
        pinned = torch.from_numpy(host_rays).pin_memory()
        out = torch.empty((host_rays.shape[0], 4), dtype=torch.float32).pin_memory()
        gb.trace(pinned, out=out)          # zero-copy: the kernel reads/writes pinned memory
        assert np.array_equal(bits(out.numpy()), bits(want))
        for zc in (0, 1, 2, 3):             # staged / zero-copy rays / zero-copy hits / both
            rt.set_param("trace.zero_copy", zc)
            try:
                out.zero_()
                gb.trace(pinned, out=out)
            finally:
                rt.set_param("trace.zero_copy", 0)
            assert np.array_equal(bits(out.numpy()), bits(want)), zc
        mixed = gb.trace(pinned)            # pinned rays -> hits of the same kind (host tensor)
        assert np.array_equal(bits(mixed.numpy()), bits(want))
        # streamed (one launch follows the upload's watermark) against staged chunks, on a ray
        # count that ends inside a segment, for every kernel flavour
        m = 1300017
        for flags in (0, rt.RT_TRACE_ANY, rt.RT_TRACE_WATERTIGHT, rt.RT_TRACE_ANY | rt.RT_TRACE_WATERTIGHT):
            ref = gb.trace(g_rays[:m], flags).cpu().numpy()
            for streamed in (1, 0):
                rt.set_param("trace.streamed", streamed)
                try:
                    out.zero_()
                    gb.trace(pinned[:m], flags, out=out[:m])
                finally:
                    rt.set_param("trace.streamed", 1)
                assert np.array_equal(bits(out[:m].numpy()), bits(ref)), (flags, streamed)
                assert not out[m:].any(), "wrote past the batch"


def test_config3_ten_million_triangles(rt, oracle):
    """Config 3 (BASELINE.json:9) at full size: the 9,979,112-triangle sphereflake.  Keys, sorted
    order, topology and node boxes against the oracle's build; a sample of the 4K primary rays
    and their AO rays against the oracle's traversal."""
    d_tris = rt.gen_scene(rt.RT_SCENE_SPHEREFLAKE, 4, 13, 2)
    tris = d_tris.cpu().numpy()
    n = tris.shape[0]
    assert n == 9979112
    for key_bits in (30, 63):
        ob = oracle.OracleBvh(tris, key_bits)
        with rt.Bvh.build(d_tris, key_bits=key_bits, keep_keys=True) as gb:
            keys, idx = gb.read_sorted()
            assert np.array_equal(keys, ob.keys) and np.array_equal(idx, ob.idx), key_bits
            left, right, parent = gb.read_topology()
            assert np.array_equal(left, ob.left) and np.array_equal(right, ob.right), key_bits
            assert np.array_equal(parent, ob.parent), key_bits
            assert np.array_equal(bits(gb.read_boxes()), bits(ob.boxes)), key_bits
            if key_bits != 30:
                continue
            cam = rt.camera.frame_box(gb.bounds, 3840, 2160)
            g_rays = rt.gen_primary_rays(cam, 3840, 2160)
            got = gb.trace(g_rays)
            pick = torch.arange(0, g_rays.shape[0], 17, device=g_rays.device)  # 488K rays
            want = ob.trace(g_rays[pick].cpu().numpy())
            assert np.array_equal(bits(got[pick].cpu().numpy()), bits(want))
            radius = 0.05 * float(np.linalg.norm(gb.bounds[3:] - gb.bounds[:3]))
            ao = rt.gen_ao_rays(d_tris, g_rays[pick], got[pick], 4, 2, radius)
            got_ao = gb.trace(ao, rt.RT_TRACE_ANY).cpu().numpy()
            want_ao = ob.trace(ao.cpu().numpy(), oracle.TRACE_ANY)
            assert np.array_equal(bits(got_ao), bits(want_ao))
            assert 0.02 < (oracle.hit_prims(got_ao) >= 0).mean() < 0.98


def test_refit_after_moving_vertices(rt, oracle):
    tris = oracle.gen_scene(oracle.SCENE_SOUP, 5000, 0, seed=3)
    moved = tris.copy()
    moved[:, [1, 5, 9]] += np.float32(0.25) * np.sin(moved[:, [0, 4, 8]] * 7).astype(np.float32)
    rays = oracle.gen_random_rays(np.array([0, 0, 0, 1, 1.3, 1], dtype=np.float32), 1, 0, 20000)
    with rt.Bvh.build(dev(tris)) as gb:
        left0, right0, parent0 = gb.read_topology()
        gb.refit(dev(moved))
        left, right, parent = gb.read_topology()
        assert np.array_equal(left, left0) and np.array_equal(parent, parent0)
        _, idx = gb.read_sorted(keys=False)
        check_boxes(gb.n_tris, left, right, gb.read_boxes(), moved, idx)
        got = gb.trace(dev(rays)).cpu().numpy()
    want = oracle.brute(moved, rays)
    assert np.array_equal(bits(got), bits(want)), "BVH traversal == all-pairs intersector, exactly"


def test_concurrent_trace_on_one_handle(rt, oracle):
    tris = oracle.gen_scene(oracle.SCENE_SOUP, 20000, 0, seed=1)
    ob = oracle.OracleBvh(tris)
    rays = [oracle.gen_random_rays(ob.bounds, 4, k * 30000, 30000) for k in range(4)]
    want = [ob.trace(r) for r in rays]
    got = [None] * 4
    with rt.Bvh.build(dev(tris)) as gb:
        def work(k):
            got[k] = gb.trace(rays[k])  # host buffers: per-thread pipeline
        ts = [threading.Thread(target=work, args=(k,)) for k in range(4)]
        [t.start() for t in ts]
        [t.join() for t in ts]
    for k in range(4):
        assert np.array_equal(bits(got[k]), bits(want[k]))


def test_launch_counter_counts(rt, oracle):
    tris = dev(oracle.gen_scene(oracle.SCENE_CORNELL))
    before = rt.kernel_launches()
    with rt.Bvh.build(tris) as gb:
        gb.trace(rt.gen_primary_rays(rt.camera.cornell_camera(64, 64), 64, 64))
    assert rt.kernel_launches() - before >= 10


# ---- knobs that must not change a bit ---------------------------------------------------
@pytest.mark.parametrize("pdl", [0, 1])
def test_build_with_and_without_dependent_launch(rt, oracle, pdl):
    """`build.pdl`: the build chain launched as programmatic dependents (default) or
    as plain stream-ordered launches; both must give the oracle's tree, also when the
    same pool memory is rebuilt back to back with different scenes."""
    rt.set_param("build.pdl", pdl)
    try:
        for rep in range(3):
            for n, key_bits in [(40001, 30), (7681, 63), (200000, 30)]:
                tris = oracle.gen_scene(oracle.SCENE_SOUP, n, 0, seed=100 + rep)
                # no stage events: nothing but kernels between the first and the last launch
                _build_and_compare(rt, oracle, tris, key_bits, timings=False)
    finally:
        rt.set_param("build.pdl", 1)


@pytest.mark.parametrize("levels", [8, 16, 24])
def test_shared_memory_stack_top(rt, oracle, levels):
    """`trace.sh_stack`: the first levels of the traversal stack in shared memory."""
    tris = oracle.gen_scene(oracle.SCENE_SOUP, 20000, 0, seed=5)
    v = tris.reshape(-1, 4)[:, :3]
    box = np.concatenate([v.min(0), v.max(0)])
    rays = oracle.gen_random_rays(box, 4, 0, 60000)
    rt.set_param("trace.sh_stack", levels)
    try:
        for flags in (0, 1):
            got, want = _trace_both(rt, oracle, tris, rays, flags)
            assert np.array_equal(bits(got), bits(want)), (levels, flags)
    finally:
        rt.set_param("trace.sh_stack", -1)


@pytest.mark.parametrize("name", ["soup20k", "flake"])
def test_reordered_trace_is_bit_identical(rt, oracle, name):
    """RT_TRACE_REORDER walks the rays in coherence-key order; every hit must land at its
    ray's own position with the oracle's bits, for closest-hit, any-hit and watertight."""
    kind, p0, p1, seed = SCENES[name]
    tris = oracle.gen_scene(kind, p0, p1, seed)
    v = tris.reshape(-1, 4)[:, :3]
    box = np.concatenate([v.min(0), v.max(0)])
    rays = oracle.gen_random_rays(box, 4, 0, 150001)  # ragged, above the 64K threshold
    rays[7, 4:7] = 0.0          # a zero direction
    rays[11, 0:3] = box[3:] * 4  # an origin far outside the scene box
    ob = oracle.OracleBvh(tris, 30)
    with rt.Bvh.build(dev(tris)) as gb:
        d_rays = dev(rays)
        for flags in (0, rt.RT_TRACE_ANY, rt.RT_TRACE_WATERTIGHT):
            want = ob.trace(rays, flags)
            plain = gb.trace(d_rays, flags).cpu().numpy()
            got = gb.trace(d_rays, flags | rt.RT_TRACE_REORDER).cpu().numpy()
            assert np.array_equal(bits(got), bits(plain)), flags
            assert np.array_equal(bits(got), bits(want)), flags
        # host buffers: the flag is accepted and ignored
        got = gb.trace(rays, rt.RT_TRACE_REORDER)
        assert np.array_equal(bits(np.asarray(got)), bits(ob.trace(rays, 0)))


def test_hybrid_63_bit_sort(rt, oracle):
    """`sort.hybrid`: 63-bit builds sort the top 30 bits with the packed passes and rank
    inside the runs of equal top bits; long runs (all centroids equal) fall back to the
    seven-pass sort.  Either way the oracle's order, tree and boxes."""
    rt.set_param("sort.hybrid", 1)
    try:
        for name in SCENES:
            kind, p0, p1, seed = SCENES[name]
            _build_and_compare(rt, oracle, oracle.gen_scene(kind, p0, p1, seed), 63, timings=False)
        for n in (2, 3, 255, 256, 257, 5121, 40001, 300000):
            _build_and_compare(rt, oracle, oracle.gen_scene(oracle.SCENE_SOUP, n, 0, seed=n), 63)
        base = oracle.gen_scene(oracle.SCENE_SOUP, 300, 0, seed=2)
        _build_and_compare(rt, oracle, np.repeat(base, 40, axis=0), 63)       # runs of 40
        _build_and_compare(rt, oracle, np.repeat(base[:7], 256, axis=0), 63)  # runs of exactly 256
        _build_and_compare(rt, oracle, np.repeat(base[:7], 257, axis=0), 63)  # 257: falls back
        _build_and_compare(rt, oracle, np.repeat(base[:1], 5000, axis=0), 63)  # one run: falls back
        # and the plain seven-pass order, which the default no longer exercises
        rt.set_param("sort.hybrid", 0)
        for name in SCENES:
            kind, p0, p1, seed = SCENES[name]
            _build_and_compare(rt, oracle, oracle.gen_scene(kind, p0, p1, seed), 63, timings=False)
        _build_and_compare(rt, oracle, oracle.gen_scene(oracle.SCENE_SOUP, 40001, 0, seed=9), 63)
    finally:
        rt.set_param("sort.hybrid", 1)
