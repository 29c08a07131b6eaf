"""Golden fixtures (tests/golden/*.npz, made by tests/golden/make_golden.py).

CPU: the oracle still reproduces every committed byte.  GPU: the CUDA path,
through the C ABI, reproduces them too — this leg needs neither the oracle's
build step nor /root/reference on the GPU box.
"""
import glob
import hashlib
import os

import numpy as np
import pytest

from conftest import bits

HERE = os.path.dirname(os.path.abspath(__file__))
FIXTURES = sorted(glob.glob(os.path.join(HERE, "golden", "*.npz")))
SCENES = {
    "cornell_96": (0, 0, 0, 0), "soup3000_k30": (1, 3000, 0, 1), "soup3000_k63": (1, 3000, 0, 1),
    "flake_d2_m3": (2, 2, 3, 0), "terrain_48": (3, 48, 16, 3),
}


def digest(a):
    return np.frombuffer(hashlib.sha256(np.ascontiguousarray(a).tobytes()).digest(), dtype=np.uint8)


def _name(path):
    return os.path.splitext(os.path.basename(path))[0]


def test_fixtures_exist():
    assert len(FIXTURES) == len(SCENES)


@pytest.mark.parametrize("path", FIXTURES, ids=_name)
def test_oracle_reproduces_golden(oracle, path):
    g = np.load(path)
    kind, p0, p1, seed = SCENES[_name(path)]
    tris = oracle.gen_scene(kind, p0, p1, seed)
    assert np.array_equal(digest(tris), g["tris_sha"])
    rays = (oracle.gen_primary_rays(g["cam"], 96, 96) if "cam" in g.files
            else oracle.gen_random_rays(g["box"], 4, 0, 4096))
    assert np.array_equal(digest(rays), g["rays_sha"])
    b = oracle.OracleBvh(tris, int(g["key_bits"]))
    for k in ("keys", "idx", "left", "right", "parent"):
        assert np.array_equal(getattr(b, k), g[k]), k
    assert np.array_equal(bits(b.bounds), bits(g["bounds"]))
    assert np.array_equal(digest(b.boxes), g["boxes_sha"])
    assert np.array_equal(bits(b.trace(rays)), bits(g["hits"]))
    assert np.array_equal(oracle.hit_prims(b.trace(rays, oracle.TRACE_ANY)), g["any_prim"])
    assert np.array_equal(bits(b.trace(rays, oracle.TRACE_WATERTIGHT)), bits(g["wt_hits"]))
    wide_boxes, wide_desc = b.collapse4(2)
    assert np.array_equal(wide_desc, g["wide_desc"])
    assert np.array_equal(digest(wide_boxes), g["wide_boxes_sha"])


@pytest.mark.gpu
@pytest.mark.parametrize("path", FIXTURES, ids=_name)
def test_gpu_reproduces_golden(rt, path):
    g = np.load(path)
    kind, p0, p1, seed = SCENES[_name(path)]
    tris = rt.gen_scene(kind, p0, p1, seed)
    assert np.array_equal(digest(tris.cpu().numpy()), g["tris_sha"])
    rays = (rt.gen_primary_rays(g["cam"], 96, 96) if "cam" in g.files
            else rt.gen_random_rays(g["box"], 4, 0, 4096))
    assert np.array_equal(digest(rays.cpu().numpy()), g["rays_sha"])
    with rt.Bvh.build(tris, key_bits=int(g["key_bits"]), keep_keys=True) as b:
        keys, idx = b.read_sorted()
        left, right, parent = b.read_topology()
        assert np.array_equal(keys, g["keys"]) and np.array_equal(idx, g["idx"])
        assert np.array_equal(left, g["left"]) and np.array_equal(right, g["right"])
        assert np.array_equal(parent, g["parent"])
        assert np.array_equal(bits(b.bounds), bits(g["bounds"]))
        boxes = b.read_boxes()
        assert np.array_equal(digest(boxes), g["boxes_sha"])
        assert np.array_equal(bits(boxes[:64]), bits(g["boxes_head"]))
        assert np.array_equal(bits(b.trace(rays).cpu().numpy()), bits(g["hits"]))
        occ = b.trace(rays, rt.RT_TRACE_ANY).cpu().numpy()
        assert np.array_equal(np.ascontiguousarray(occ[:, 3]).view(np.int32), g["any_prim"])
        wt = b.trace(rays, rt.RT_TRACE_WATERTIGHT).cpu().numpy()
        assert np.array_equal(bits(wt), bits(g["wt_hits"]))
        # A5w: the collapsed tree, then the same hits through it
        from test_wide import gpu_preorder
        b.widen()
        nodes, root = b.read_wide()
        wide_boxes, wide_desc = gpu_preorder(nodes, root, b.n_tris)
        assert np.array_equal(wide_desc, g["wide_desc"])
        assert np.array_equal(digest(wide_boxes), g["wide_boxes_sha"])
        assert np.array_equal(bits(b.trace(rays).cpu().numpy()), bits(g["hits"]))
        assert np.array_equal(bits(b.trace(rays, rt.RT_TRACE_WATERTIGHT).cpu().numpy()), bits(g["wt_hits"]))
