"""A5w, second wide layout: the compressed 8-wide BVH (csrc/wide8.cu).

CPU: the oracle's collapse8 is a well-formed tree whose quantised boxes contain what they
stand for.  GPU: the device's tree, renumbered breadth first, equals the oracle's byte for
byte; tracing through it returns the oracle's (binary-tree) hits bit for bit: closest,
any-hit, watertight, reordered, zero direction components, after a refit, after a load.
"""
import numpy as np
import pytest

from conftest import bits
from test_host_logic import zero_direction_cases

NODE8 = np.dtype([("p", np.float32, 3), ("e", np.uint8, 3), ("imask", np.uint8),
                  ("child_base", np.uint32), ("tri_base", np.uint32), ("meta", np.uint8, 8),
                  ("qlo", np.uint8, (3, 8)), ("qhi", np.uint8, (3, 8))])
SCENES = {"cornell": (0, 0, 0, 0), "soup20k": (1, 20000, 0, 1), "flake": (2, 2, 4, 0), "terrain": (3, 64, 16, 3)}


def check_wide8(nodes, tris8, n):
    """Every node and every triangle reached exactly once from node 0; internal children stored
    in slot order; leaf children's triangles inside their dequantised box."""
    nd = nodes.view(NODE8).reshape(-1)
    m = nd.shape[0]
    seen_node = np.zeros(m, dtype=bool)
    seen_tri = np.zeros(n, dtype=bool)
    stack = [0]
    while stack:
        i = stack.pop()
        assert not seen_node[i]
        seen_node[i] = True
        node = nd[i]
        cell = np.ldexp(1.0, node["e"].astype(np.int64) - 127) * (node["e"] != 0)
        p = node["p"].astype(np.float64)
        imask = int(node["imask"])
        kids = 0
        for s in range(8):
            mt = int(node["meta"][s])
            if mt == 0:
                assert not (imask >> s) & 1
                continue
            kids += 1
            lo = p + node["qlo"][:, s] * cell
            hi = p + node["qhi"][:, s] * cell
            if (mt & 0x18) == 0x18 and (mt >> 5) == 1:
                assert (mt & 7) == s and (imask >> s) & 1
                child = int(node["child_base"]) + bin(imask & ((1 << s) - 1)).count("1")
                assert (nd[child]["p"].astype(np.float64) >= lo).all(), "child's min corner inside its slot box"
                stack.append(child)
            else:
                assert not (imask >> s) & 1
                cnt, off = bin(mt >> 5).count("1"), mt & 31
                assert 1 <= cnt <= 3 and (mt >> 5) in (1, 3, 7) and off + cnt <= 24
                for t in range(cnt):
                    j = int(node["tri_base"]) + off + t
                    assert not seen_tri[j]
                    seen_tri[j] = True
                    v = tris8[j].reshape(3, 4)[:, :3].astype(np.float64)
                    assert (v.min(0) >= lo).all() and (v.max(0) <= hi).all(), "leaf box contains its triangles"
        assert kids >= 2
    assert seen_node.all() and seen_tri.all()


@pytest.mark.parametrize("key_bits", [30, 63])
@pytest.mark.parametrize("name", list(SCENES))
def test_oracle_collapse8_is_well_formed(oracle, name, key_bits):
    kind, p0, p1, seed = SCENES[name]
    tris = oracle.gen_scene(kind, p0, p1, seed)
    if name == "soup20k":
        tris = tris[:6000]
    ob = oracle.OracleBvh(tris, key_bits)
    nodes, tris8 = ob.collapse8()
    check_wide8(nodes, tris8, ob.n)
    prim = tris8[:, 3].copy().view(np.uint32)
    assert np.array_equal(np.sort(prim), np.arange(ob.n, dtype=np.uint32))
    assert np.array_equal(bits(tris8[:, [0, 1, 2, 4, 5, 6, 8, 9, 10]]), bits(tris[prim][:, [0, 1, 2, 4, 5, 6, 8, 9, 10]]))
    cn, ct = oracle.canon8(nodes, tris8)
    assert np.array_equal(cn, nodes) and np.array_equal(bits(ct), bits(tris8)), "collapse8 is canonical"
    fill = (nodes.view(NODE8).reshape(-1)["meta"] != 0).sum(1).mean()
    assert fill > 5.0, fill
    assert nodes.shape[0] * 80 < 0.3 * (ob.n - 1) * 64, "less than a third of the binary node array"
    ob.close()


def test_oracle_collapse8_small(oracle):
    base = oracle.gen_scene(oracle.SCENE_SOUP, 64, 0, seed=3)
    for n in (1, 2, 3):
        assert oracle.OracleBvh(base[:n]).collapse8()[0].shape[0] == 0
    for n in (4, 5, 9, 25, 64):
        ob = oracle.OracleBvh(base[:n])
        nodes, tris8 = ob.collapse8()
        check_wide8(nodes, tris8, n)
    dup = np.repeat(base[:3], 50, axis=0)          # 150 triangles in three Morton cells
    ob = oracle.OracleBvh(dup)
    check_wide8(*ob.collapse8(), 150)


# ---- GPU -------------------------------------------------------------------------------------
gpu = pytest.mark.gpu


def dev(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


@gpu
@pytest.mark.parametrize("key_bits", [30, 63])
@pytest.mark.parametrize("name", list(SCENES))
def test_device_tree_equals_the_oracles(rt, oracle, name, key_bits):
    kind, p0, p1, seed = SCENES[name]
    tris = oracle.gen_scene(kind, p0, p1, seed)
    ob = oracle.OracleBvh(tris, key_bits)
    want_nodes, want_tris = ob.collapse8()
    with rt.Bvh.build(dev(tris), key_bits=key_bits, wide8=True) as gb:
        nodes, tris8 = gb.read_wide8()
    assert nodes.shape[0] == want_nodes.shape[0]
    got_nodes, got_tris = oracle.canon8(nodes, tris8)
    assert np.array_equal(bits(got_tris), bits(want_tris)), "triangles in wide-node order"
    assert np.array_equal(got_nodes, want_nodes), "wide nodes (breadth-first numbering)"
    check_wide8(nodes.view(np.uint8).reshape(-1, 80), tris8, ob.n)
    ob.close()


@gpu
@pytest.mark.parametrize("name", list(SCENES))
def test_wide8_hits_are_bit_identical(rt, oracle, name):
    kind, p0, p1, seed = SCENES[name]
    tris = oracle.gen_scene(kind, p0, p1, seed)
    v = tris.reshape(-1, 4)[:, :3]
    box = np.concatenate([v.min(0), v.max(0)])
    rays = oracle.gen_random_rays(box, 4, 0, 150001)
    cam = rt.camera.frame_box(box, 160, 120)
    rays = np.concatenate([rays, oracle.gen_primary_rays(cam, 160, 120)])
    rays[7, 4:7] = 0.0
    rays[11, 0:3] = box[3:] * 4
    d_rays = dev(rays)
    for key_bits in (30, 63):
        ob = oracle.OracleBvh(tris, key_bits)
        with rt.Bvh.build(dev(tris), key_bits=key_bits, wide8=True) as gb:
            for flags in (0, rt.RT_TRACE_ANY, rt.RT_TRACE_WATERTIGHT, rt.RT_TRACE_ANY | rt.RT_TRACE_WATERTIGHT):
                want = ob.trace(rays, flags)
                got = gb.trace(d_rays, flags).cpu().numpy()
                assert np.array_equal(bits(got), bits(want)), (key_bits, flags)
                got = gb.trace(d_rays, flags | rt.RT_TRACE_REORDER).cpu().numpy()
                assert np.array_equal(bits(got), bits(want)), (key_bits, flags, "reorder")
            # host buffers (staged pipeline) and the same handle on the binary layout
            got = gb.trace(rays)
            assert np.array_equal(bits(np.asarray(got)), bits(ob.trace(rays)))
            rt.set_param("trace.wide", 0)
            try:
                got2 = gb.trace(d_rays).cpu().numpy()
            finally:
                rt.set_param("trace.wide", 1)
            assert np.array_equal(bits(got2), bits(ob.trace(rays)))
            hits, n_nodes, n_tests = gb.trace_stats(d_rays)
            assert np.array_equal(bits(hits.cpu().numpy()), bits(ob.trace(rays)))
            _, o_nodes, o_tests = (None,) + ob.trace(rays, counters=True)[1]
            assert 0 < n_nodes < o_nodes, "fewer node fetches than the binary tree"
            assert n_tests > 0
        ob.close()


@gpu
def test_wide8_zero_direction_rays_match_brute_force(rt, oracle):
    tris, rays = zero_direction_cases()
    with rt.Bvh.build(dev(tris), wide8=True) as gb:
        for flags in (rt.RT_TRACE_WATERTIGHT, 0, rt.RT_TRACE_ANY | rt.RT_TRACE_WATERTIGHT):
            got = gb.trace(dev(rays), flags).cpu().numpy()
            assert np.array_equal(bits(got), bits(oracle.brute(tris, rays, flags))), flags


@gpu
def test_wide8_edge_cases_refit_and_load(rt, oracle, tmp_path):
    base = oracle.gen_scene(oracle.SCENE_SOUP, 5000, 0, seed=3)
    ray = oracle.gen_random_rays(np.array([0, 0, 0, 1, 1, 1], dtype=np.float32), 1, 0, 20000)
    # too small for the layout: the flag is accepted, the binary tree answers
    for n in (0, 1, 2, 3):
        with rt.Bvh.build(base[:n], wide8=True) as gb:
            want = oracle.OracleBvh(base[:n]).trace(ray) if n else None
            got = gb.trace(ray)
            if n:
                assert np.array_equal(bits(got), bits(want))
            else:
                assert (oracle.hit_prims(got) == -1).all()
            with pytest.raises(rt.RtError):
                gb.read_wide8()
    for n in (4, 5, 9, 33):
        with rt.Bvh.build(base[:n], wide8=True) as gb:
            nodes, tris8 = gb.read_wide8()
            check_wide8(nodes.view(np.uint8).reshape(-1, 80), tris8, n)
            assert np.array_equal(bits(gb.trace(ray)), bits(oracle.OracleBvh(base[:n]).trace(ray)))
    dup = np.repeat(base[:3], 50, axis=0)
    with rt.Bvh.build(dup, key_bits=63, wide8=True) as gb:
        assert np.array_equal(bits(gb.trace(ray)), bits(oracle.OracleBvh(dup, 63).trace(ray)))
    # refit: moved vertices, same topology, the layout is rebuilt
    moved = base.copy()
    moved[:, [1, 5, 9]] += np.float32(0.25) * np.sin(moved[:, [0, 4, 8]] * 7).astype(np.float32)
    with rt.Bvh.build(dev(base), wide8=True) as gb:
        gb.refit(dev(moved))
        got = gb.trace(dev(ray)).cpu().numpy()
        rt.set_param("trace.wide", 0)
        try:
            plain = gb.trace(dev(ray)).cpu().numpy()
        finally:
            rt.set_param("trace.wide", 1)
        assert np.array_equal(bits(got), bits(plain)), "8-wide == binary after a refit"
        want = oracle.brute(moved, ray)
        assert np.array_equal(bits(got), bits(want)), "BVH traversal == all-pairs intersector, exactly"
        # a saved handle has no wide layout; bvh_widen8 gives it one
        path = tmp_path / "w8.bvh"
        gb.save(str(path))
    with rt.Bvh.load(str(path)) as lb:
        with pytest.raises(rt.RtError):
            lb.read_wide8()
        lb.widen8()
        nodes, tris8 = lb.read_wide8()
        check_wide8(nodes.view(np.uint8).reshape(-1, 80), tris8, lb.n_tris)
        assert np.array_equal(bits(lb.trace(dev(ray)).cpu().numpy()), bits(got))
