"""Seeded fuzzing of the acceptance rule: every layout and every traversal returns
the bytes of the all-pairs intersector, whatever the scene and the rays.

Scenes mix small random triangles with the things that break traversals: zero-area triangles,
exact duplicates, coplanar axis-aligned quads on shared planes, slivers, one huge triangle over
everything; shifted and scaled by up to 10^4.  Rays mix random ones with exact-zero direction
components, origins exactly on vertices and planes, finite reach and positive tmin.

CPU: the oracle's binary tree (both key widths) against its brute force.  GPU: the binary tree,
the compact nodes, both wide layouts, reordered launches, any-hit, watertight.
"""
import numpy as np
import pytest

from conftest import bits

SEEDS = list(range(12))
GPU_SEEDS = list(range(36))


def fuzz_scene(seed):
    rng = np.random.default_rng(1000 + seed)
    n = int(rng.integers(40, 1500))
    c = rng.random((n, 1, 3))
    size = 10.0 ** rng.uniform(-3, -0.7, (n, 1, 1))
    v = c + (rng.random((n, 3, 3)) - 0.5) * size
    k = n // 10
    v[:k, 1] = v[:k, 0]                                     # zero-area: two equal vertices
    v[k:2 * k, 2] = 0.5 * (v[k:2 * k, 0] + v[k:2 * k, 1])   # zero-area: collinear
    v[2 * k:3 * k] = v[3 * k:4 * k]                         # exact duplicates
    g = np.round(v[4 * k:6 * k] * 8) / 8                    # coplanar, axis-aligned, on shared grid planes
    g[:, :, int(rng.integers(0, 3))] = np.round(rng.random() * 4) / 4
    v[4 * k:6 * k] = g
    v[6 * k:7 * k, 2] = v[6 * k:7 * k, 0] + (v[6 * k:7 * k, 1] - v[6 * k:7 * k, 0]) * 1.0000001 + 1e-7   # slivers
    v[-1] = [[-3, -3, 0.37], [5, -3, 0.37], [1, 6, 0.37]]   # one triangle across the whole scene
    scale = [1.0, 1.0e-3, 1.0e4][seed % 3]
    shift = [0.0, 100.0, -1.0e4][(seed // 3) % 3] * scale
    v = v * scale + shift
    tris = np.zeros((n, 12), dtype=np.float32)
    tris.reshape(n, 3, 4)[:, :, :3] = v.astype(np.float32)
    return rng, tris


def fuzz_rays(rng, tris, m=6000):
    v = tris.reshape(-1, 4)[:, :3].astype(np.float64)
    lo, hi = v.min(0), v.max(0)
    ext = np.maximum(hi - lo, 1e-30)
    o = lo + (rng.random((m, 3)) * 1.6 - 0.3) * ext
    d = rng.normal(size=(m, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    z = rng.random((m, 3)) < 0.15
    z[z.all(1), 0] = False
    d[z] = rng.choice([0.0, -0.0], size=int(z.sum()))       # exact zeros of either sign
    on_vertex = rng.random(m) < 0.1
    o[on_vertex] = v[rng.integers(0, len(v), int(on_vertex.sum()))]
    toward = rng.random(m) < 0.3                             # aimed at a vertex: edge and vertex hits
    tgt = v[rng.integers(0, len(v), m)]
    aim = tgt - o
    ln = np.linalg.norm(aim, axis=1, keepdims=True)
    ok = toward & (ln[:, 0] > 0)
    d[ok] = (aim / np.where(ln > 0, ln, 1))[ok]
    rays = np.zeros((m, 8), dtype=np.float32)
    rays[:, :3] = o
    rays[:, 4:7] = d
    rays[:, 3] = np.where(rng.random(m) < 0.2, rng.random(m) * 0.2 * np.linalg.norm(ext), 0.0)   # tmin
    rays[:, 7] = np.where(rng.random(m) < 0.3, rng.random(m) * np.linalg.norm(ext), np.inf)      # reach
    return rays


@pytest.mark.parametrize("seed", SEEDS)
def test_oracle_trees_equal_brute_force(oracle, seed):
    rng, tris = fuzz_scene(seed)
    rays = fuzz_rays(rng, tris, 2500)
    for flags in (oracle.TRACE_CLOSEST, oracle.TRACE_ANY, oracle.TRACE_WATERTIGHT):
        want = oracle.brute(tris, rays, flags)
        for key_bits in (30, 63):
            ob = oracle.OracleBvh(tris, key_bits)
            assert np.array_equal(bits(ob.trace(rays, flags)), bits(want)), (seed, flags, key_bits)
            ob.close()
    assert (oracle.hit_prims(oracle.brute(tris, rays)) >= 0).mean() > 0.05


# ---- GPU -------------------------------------------------------------------------------------
def dev(a):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


@pytest.mark.gpu
@pytest.mark.parametrize("seed", GPU_SEEDS)
def test_every_layout_equals_brute_force(rt, oracle, seed):
    rng, tris = fuzz_scene(seed)
    rays = fuzz_rays(rng, tris)
    d_rays = dev(rays)
    want = {f: oracle.brute(tris, rays, f) for f in (0, oracle.TRACE_ANY, oracle.TRACE_WATERTIGHT,
                                                     oracle.TRACE_ANY | oracle.TRACE_WATERTIGHT)}
    for key_bits in (30, 63):
        for layout in ({}, {"compact": True}, {"wide8": True}, {"wide": True}):
            with rt.Bvh.build(dev(tris), key_bits=key_bits, **layout) as gb:
                for flags, w in want.items():
                    got = gb.trace(d_rays, flags).cpu().numpy()
                    assert np.array_equal(bits(got), bits(w)), (seed, key_bits, layout, flags)
                got = gb.trace(d_rays, rt.RT_TRACE_REORDER).cpu().numpy()
                assert np.array_equal(bits(got), bits(want[0])), (seed, key_bits, layout, "reorder")
    got = rt.Bvh.build(tris).trace(rays)                    # host buffers
    assert np.array_equal(bits(np.asarray(got)), bits(want[0]))


@pytest.mark.gpu
@pytest.mark.parametrize("seed", SEEDS)
def test_degenerate_scenes_build_the_oracles_tree(rt, oracle, seed):
    """keys, order (duplicates and equal centroids: ties by index), topology and boxes of the
    fuzz scenes, both key widths, byte for byte"""
    _, tris = fuzz_scene(seed)
    for key_bits in (30, 63):
        ob = oracle.OracleBvh(tris, key_bits)
        with rt.Bvh.build(dev(tris), key_bits=key_bits, keep_keys=True) as gb:
            keys, idx = gb.read_sorted()
            left, right, parent = gb.read_topology()
            boxes = gb.read_boxes()
            assert np.array_equal(bits(gb.bounds), bits(ob.bounds))
        assert np.array_equal(keys, ob.keys) and np.array_equal(idx, ob.idx), (seed, key_bits)
        assert np.array_equal(left, ob.left) and np.array_equal(right, ob.right) and np.array_equal(parent, ob.parent)
        assert np.array_equal(bits(boxes), bits(ob.boxes)), (seed, key_bits)
        ob.close()
