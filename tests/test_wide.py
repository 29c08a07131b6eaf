"""A5w: the 4-wide BVH.

CPU: the oracle's collapse obeys the structural invariants (every sorted triangle in
exactly one reachable leaf, leaves of at most 4 triangles, child boxes = the binary
tree's boxes).  GPU: the CUDA collapse equals the oracle's wide tree array for array
(pre-order), and tracing through the wide layout returns bit-identical hits to the
oracle and to the binary layout, for closest-hit and any-hit rays.
"""
import numpy as np
import pytest

from conftest import bits

SCENES = {
    "cornell": (0, 0, 0, 0),
    "soup20k": (1, 20000, 0, 1),
    "flake": (2, 2, 4, 0),
    "terrain": (3, 64, 16, 3),
}


def check_wide_invariants(ob, boxes, desc, leaf_max=4):
    n = ob.n
    m = desc.shape[0]
    if n <= leaf_max:
        assert m == 0
        return
    kind = desc[:, :, 0]
    assert (kind[:, :2] != 0).all(), "a wide node has at least two children"
    assert ((kind == 0)[:, :-1] <= (kind == 0)[:, 1:]).all(), "empty slots come last"
    leaves = desc[kind == 1]
    assert (leaves[:, 2] >= 1).all() and (leaves[:, 2] <= leaf_max).all()
    cover = np.zeros(n, dtype=np.int32)
    for f, c in leaves[:, 1:]:
        cover[f:f + c] += 1
    assert (cover == 1).all(), "every sorted triangle lies in exactly one leaf"
    children = desc[kind == 2][:, 1]
    assert np.array_equal(np.sort(children), np.arange(1, m)), "pre-order: each node one parent"
    # a leaf's box bounds its triangles exactly (it is the binary subtree's box)
    t = ob._tris.reshape(n, 3, 4)[:, :, :3][ob.idx]
    lo, hi = t.min(axis=1), t.max(axis=1)
    lb = boxes[kind == 1]
    for (f, c), b in list(zip(leaves[:, 1:], lb))[:2000]:
        assert np.array_equal(b[:3], lo[f:f + c].min(axis=0))
        assert np.array_equal(b[3:], hi[f:f + c].max(axis=0))


@pytest.mark.parametrize("name", list(SCENES))
def test_oracle_collapse_invariants(oracle, name):
    kind, p0, p1, seed = SCENES[name]
    ob = oracle.OracleBvh(oracle.gen_scene(kind, p0, p1, seed))
    for leaf_max in (1, 2, 4):
        boxes, desc = ob.collapse4(leaf_max)
        check_wide_invariants(ob, boxes, desc, leaf_max)
        assert desc.shape[0] < ob.n / 2, "fewer wide nodes than binary nodes"


def test_oracle_collapse_tiny(oracle):
    tri = np.array([[0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0]], dtype=np.float32)
    for n in (1, 2, 4, 5, 9):
        tris = np.concatenate([tri + np.float32(k) for k in range(n)])
        ob = oracle.OracleBvh(tris)
        for leaf_max in (1, 2, 4):
            boxes, desc = ob.collapse4(leaf_max)
            check_wide_invariants(ob, boxes, desc, leaf_max)


# ---- GPU ----------------------------------------------------------------------------------
LEAF_DEFAULT = 2  # kWideLeafDefault (csrc/kernels.h)


def gpu_preorder(nodes, root_ref, n, leaf_max=LEAF_DEFAULT):
    """The reachable part of bvh_read_wide in the oracle's pre-order form."""
    from gpu_raytracer_b200.api import WIDE_EMPTY, WIDE_FIRST_MASK, WIDE_LEAF_BIT
    boxes, desc = [], []
    if n <= leaf_max:
        return np.zeros((0, 4, 6), np.float32), np.zeros((0, 4, 3), np.uint32)

    def visit(g):
        me = len(desc)
        desc.append(np.zeros((4, 3), np.uint32))
        bx = np.zeros((4, 6), np.float32)
        boxes.append(bx)
        nd = nodes[g]
        for k in range(4):
            ref = int(nd["ref"][k])
            if ref == WIDE_EMPTY:
                continue
            bx[k, :3] = nd["lo"][:, k]
            bx[k, 3:] = nd["hi"][:, k]
            if ref & WIDE_LEAF_BIT:
                desc[me][k] = (1, ref & WIDE_FIRST_MASK, ((ref >> 29) & 3) + 1)
            else:
                desc[me][k] = (2, visit(ref), 0)
        return me

    visit(root_ref)
    return np.stack(boxes), np.stack(desc)


@pytest.mark.gpu
@pytest.mark.parametrize("key_bits,leaf_max", [(30, 2), (63, 2), (30, 1), (30, 4), (63, 3)])
@pytest.mark.parametrize("name", list(SCENES))
def test_gpu_collapse_matches_oracle(rt, oracle, name, key_bits, leaf_max):
    import torch
    kind, p0, p1, seed = SCENES[name]
    tris = oracle.gen_scene(kind, p0, p1, seed)
    ob = oracle.OracleBvh(tris, key_bits)
    want_boxes, want_desc = ob.collapse4(leaf_max)
    rays = oracle.gen_random_rays(ob.bounds, 4, 0, 20000)
    rt.set_param("wide.leaf_max", leaf_max)
    try:
        with rt.Bvh.build(torch.from_numpy(tris).cuda(), key_bits=key_bits, wide=True, timings=True) as gb:
            nodes, root = gb.read_wide()
            got_boxes, got_desc = gpu_preorder(nodes, root, gb.n_tris, leaf_max)
            assert gb.build_times()["wide"] > 0
            hits = gb.trace(rays)
    finally:
        rt.set_param("wide.leaf_max", LEAF_DEFAULT)
    assert np.array_equal(got_desc, want_desc), "A5w wide topology"
    assert np.array_equal(bits(got_boxes), bits(want_boxes)), "A5w wide boxes"
    check_wide_invariants(ob, got_boxes, got_desc, leaf_max)
    assert np.array_equal(bits(hits), bits(ob.trace(rays)))


@pytest.mark.gpu
@pytest.mark.parametrize("name", list(SCENES))
def test_wide_hits_are_bit_identical(rt, oracle, name):
    import torch
    kind, p0, p1, seed = SCENES[name]
    tris = oracle.gen_scene(kind, p0, p1, seed)
    ob = oracle.OracleBvh(tris)
    rays = oracle.gen_random_rays(ob.bounds, 4, 0, 100000)
    d_rays = torch.from_numpy(rays).cuda()
    with rt.Bvh.build(torch.from_numpy(tris).cuda(), wide=True) as gb:
        for flags in (0, 1):
            want = ob.trace(rays, flags)
            got = gb.trace(d_rays, flags).cpu().numpy()
            assert np.array_equal(bits(got), bits(want)), f"wide vs oracle, flags={flags}"
            rt.set_param("trace.wide", 0)
            try:
                got2 = gb.trace(d_rays, flags).cpu().numpy()
            finally:
                rt.set_param("trace.wide", 1)
            assert np.array_equal(bits(got2), bits(want)), "binary layout of the same handle"
        # instrumented kernel: same hits; fewer node fetches than the binary walk
        s_hits, nn, nt = gb.trace_stats(d_rays)
        assert np.array_equal(bits(s_hits.cpu().numpy()), bits(ob.trace(rays)))
        nn2, _ = ob.trace(rays, counters=True)[1]
        assert 0 < nn < nn2
        # host buffers go through the staged pipeline
        assert np.array_equal(bits(gb.trace(rays)), bits(ob.trace(rays)))


@pytest.mark.gpu
def test_wide_edge_cases_and_refit(rt, oracle):
    import torch
    ray = np.array([[0.25, 0.25, -1, 0, 0, 0, 1, np.inf], [5, 5, -1, 0, 0, 0, 1, np.inf]], dtype=np.float32)
    tri = np.array([[0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0]], dtype=np.float32)
    with rt.Bvh.build(np.zeros((0, 12), dtype=np.float32), wide=True) as gb:
        assert oracle.hit_prims(gb.trace(ray)).tolist() == [-1, -1]
    for n in (1, 2, 3, 4, 5, 6, 9, 33):
        tris = np.concatenate([tri + np.float32(0.5 * k) * np.float32([0, 0, 1, 0] * 3) for k in range(n)])
        ob = oracle.OracleBvh(tris)
        with rt.Bvh.build(tris, wide=True) as gb:
            assert np.array_equal(bits(gb.trace(ray)), bits(ob.trace(ray))), n
            nodes, root = gb.read_wide()
            gb_boxes, gb_desc = gpu_preorder(nodes, root, n)
            wb, wd = ob.collapse4(LEAF_DEFAULT)
            assert np.array_equal(gb_desc, wd) and np.array_equal(bits(gb_boxes), bits(wb)), n
    # widen after the build, then refit with moved vertices
    tris = oracle.gen_scene(oracle.SCENE_SOUP, 5000, 0, seed=3)
    moved = tris.copy()
    moved[:, [1, 5, 9]] += np.float32(0.25)
    rays = oracle.gen_random_rays(np.array([0, 0, 0, 1, 1.3, 1], dtype=np.float32), 1, 0, 20000)
    with rt.Bvh.build(torch.from_numpy(tris).cuda()) as gb:
        base = gb.trace(rays)
        gb.widen()
        assert np.array_equal(bits(gb.trace(rays)), bits(base))
        gb.refit(torch.from_numpy(moved).cuda())
        got = gb.trace(rays)
        rt.set_param("trace.wide", 0)
        try:
            got2 = gb.trace(rays)
        finally:
            rt.set_param("trace.wide", 1)
        assert np.array_equal(bits(got), bits(got2))
    want = oracle.brute(moved, rays)
    assert np.array_equal(bits(got), bits(want)), "BVH traversal == all-pairs intersector, exactly"
