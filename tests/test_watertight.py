"""A7 option: the watertight (Woop-Benthin-Wald) triangle test, RT_TRACE_WATERTIGHT.

CPU: the oracle's watertight traversal equals its watertight brute force exactly (no
BVH-vs-brute discrepancies at shared edges, unlike Moeller-Trumbore), closes the
cracks Moeller-Trumbore leaves on a tessellated closed surface, and agrees with
Moeller-Trumbore to 1e-5 away from edges.  GPU: bit-identical to the oracle through
both node layouts, closest-hit and any-hit."""
import numpy as np
import pytest

from conftest import bits

WT = 2


def grid_mesh(n=24):
    """A flat n x n quad mesh in z = 1 (2 n^2 triangles sharing every interior edge)."""
    xs = np.linspace(0, 1, n + 1, dtype=np.float32)
    tris = []
    for i in range(n):
        for j in range(n):
            a, b = (xs[i], xs[j], 1), (xs[i + 1], xs[j], 1)
            c, d = (xs[i + 1], xs[j + 1], 1), (xs[i], xs[j + 1], 1)
            tris.append([*a, 0, *b, 0, *c, 0])
            tris.append([*a, 0, *c, 0, *d, 0])
    return np.array(tris, dtype=np.float32)


def edge_rays(n=24, per=400, seed=5):
    """Rays aimed exactly at mesh vertices and at points on mesh edges."""
    rng = np.random.default_rng(seed)
    xs = np.linspace(0, 1, n + 1, dtype=np.float32)
    tgt = []
    for _ in range(per):
        i, j = rng.integers(1, n, 2)
        tgt.append((xs[i], xs[j]))                                   # a vertex
        tgt.append((xs[i], np.float32(rng.uniform(0.05, 0.95))))     # a vertical edge
        tgt.append((np.float32(rng.uniform(0.05, 0.95)), xs[j]))     # a horizontal edge
        u = np.float32(rng.uniform(0.1, 0.9))
        tgt.append((xs[i] + u * (xs[i + 1] - xs[i]), xs[j] + u * (xs[j + 1] - xs[j])))  # near a diagonal
    tgt = np.array(tgt, dtype=np.float32)
    org = np.array([0.37, 0.41, -1.0], dtype=np.float32)
    d = np.concatenate([tgt, np.ones((len(tgt), 1), np.float32)], axis=1) - org
    rays = np.zeros((len(tgt), 8), dtype=np.float32)
    rays[:, :3] = org
    rays[:, 4:7] = d
    rays[:, 7] = np.inf
    return rays


def test_oracle_watertight_has_no_cracks(oracle):
    tris, rays = grid_mesh(), edge_rays()
    ob = oracle.OracleBvh(tris)
    wt = ob.trace(rays, WT)
    assert (oracle.hit_prims(wt) >= 0).all(), "every ray aimed inside the mesh hits it"
    assert np.array_equal(bits(wt), bits(oracle.brute(tris, rays, WT))), "BVH == brute force, exactly"
    assert np.allclose(wt[:, 0], np.linalg.norm(rays[:, 4:7], axis=1) * 0 + 1.0, rtol=1e-5)  # t = 1 (dz = 2 -> z = 1)


def test_oracle_watertight_agrees_with_moeller_trumbore(oracle):
    tris = oracle.gen_scene(oracle.SCENE_SPHEREFLAKE, 2, 4)
    ob = oracle.OracleBvh(tris)
    rays = oracle.gen_random_rays(ob.bounds, 4, 0, 50000)
    mt, wt = ob.trace(rays), ob.trace(rays, WT)
    pm, pw = oracle.hit_prims(mt), oracle.hit_prims(wt)
    same = pm == pw
    assert same.mean() > 0.999
    m = same & (pm >= 0)
    assert np.allclose(mt[m, :3], wt[m, :3], rtol=1e-4, atol=1e-5)
    # any-hit
    occ_m, occ_w = ob.trace(rays, 1), ob.trace(rays, 1 | WT)
    assert (oracle.hit_prims(occ_m) == oracle.hit_prims(occ_w)).mean() > 0.999


@pytest.mark.gpu
@pytest.mark.parametrize("wide", [False, True])
def test_gpu_watertight_matches_oracle(rt, oracle, wide):
    import torch
    for tris, rays in [(grid_mesh(), edge_rays()),
                       (oracle.gen_scene(oracle.SCENE_SPHEREFLAKE, 2, 4), None),
                       (oracle.gen_scene(oracle.SCENE_SOUP, 20000, 0, 1), None),
                       (oracle.gen_scene(oracle.SCENE_CORNELL), None)]:
        ob = oracle.OracleBvh(tris)
        if rays is None:
            rays = oracle.gen_random_rays(ob.bounds, 4, 0, 100000)
        with rt.Bvh.build(torch.from_numpy(tris).cuda(), wide=wide) as gb:
            for flags in (WT, WT | 1):
                want = ob.trace(rays, flags)
                got = gb.trace(torch.from_numpy(rays).cuda(), flags).cpu().numpy()
                assert np.array_equal(bits(got), bits(want)), (flags, wide)
            assert np.array_equal(bits(gb.trace(rays, WT)), bits(ob.trace(rays, WT))), "host buffers"
            for variant in (0, 1):  # plain while-while kernel too
                rt.set_param("trace.variant", variant)
                try:
                    got = gb.trace(rays, WT)
                finally:
                    rt.set_param("trace.variant", 1)
                assert np.array_equal(bits(got), bits(ob.trace(rays, WT)))
