"""scene files and BVH files.

The OBJ and binary triangle readers are host code inside the product library, so
they are checked on CPU (no compute call).  Saving and loading a built BVH needs
the device and is a GPU test.
"""
import numpy as np
import pytest

from conftest import bits

OBJ = """# a quad, a triangle and a pentagon
o thing
v 0 0 0
v 1 0 0
v 1 1 0
v 0 1 0
vt 0.5 0.5
vn 0 0 1
f 1 2 3 4
v 0 0 1
v 1 0 1
v 0.5 1 1e0
f -3/1/1 -2/1/1 -1//1
g other
f 1/1 2/1 3/1 7 5
"""


def test_obj_loader(rt, tmp_path):
    p = tmp_path / "scene.obj"
    p.write_text(OBJ)
    tris = rt.load_obj(str(p))
    assert tris.shape == (2 + 1 + 3, 12) and tris.dtype == np.float32
    v = {1: (0, 0, 0), 2: (1, 0, 0), 3: (1, 1, 0), 4: (0, 1, 0), 5: (0, 0, 1), 6: (1, 0, 1), 7: (0.5, 1, 1)}
    want = [(1, 2, 3), (1, 3, 4), (5, 6, 7), (1, 2, 3), (1, 3, 7), (1, 7, 5)]
    for t, ids in zip(tris, want):
        got = t.reshape(3, 4)
        assert np.array_equal(got[:, :3], np.array([v[i] for i in ids], dtype=np.float32))
        assert (got[:, 3] == 0).all()


@pytest.mark.parametrize("text,msg", [
    ("v 0 0 0\nv 1 0 0\nv 0 1 0\nf 1 2 4\n", "out of range"),
    ("v 0 0\n", "malformed vertex"),
    ("v 0 0 0\nv 1 0 0\nf 1 2\n", "face with 2"),
])
def test_obj_loader_errors(rt, tmp_path, text, msg):
    p = tmp_path / "bad.obj"
    p.write_text(text)
    with pytest.raises(rt.RtError) as e:
        rt.load_obj(str(p))
    assert msg in str(e.value)
    with pytest.raises(rt.RtError):
        rt.load_obj(str(tmp_path / "missing.obj"))


def test_triangle_file_round_trip(rt, oracle, tmp_path):
    tris = oracle.gen_scene(oracle.SCENE_SOUP, 1234, 0, seed=7)
    p = tmp_path / "soup.tri"
    rt.save_tris(str(p), tris)
    back = rt.load_tris(str(p))
    assert np.array_equal(bits(back), bits(tris))
    rt.save_tris(str(p), np.zeros((0, 12), np.float32))
    assert rt.load_tris(str(p)).shape == (0, 12)
    (tmp_path / "junk.tri").write_bytes(b"not a triangle file at all")
    with pytest.raises(rt.RtError):
        rt.load_tris(str(tmp_path / "junk.tri"))
    p.write_bytes(p.read_bytes()[:8] + np.uint64(5).tobytes() + b"\0" * 48)
    with pytest.raises(rt.RtError) as e:
        rt.load_tris(str(p))
    assert "truncated" in str(e.value)


@pytest.mark.gpu
def test_obj_scene_traces_like_the_oracle(rt, oracle, tmp_path):
    """OBJ text -> rt_load_obj -> build_bvh -> trace_rays == oracle on the same triangles."""
    tris = oracle.gen_scene(oracle.SCENE_SPHEREFLAKE, 1, 3)
    lines = []
    for t in tris.reshape(-1, 3, 4):
        for v in t:
            lines.append("v %s %s %s" % tuple(repr(float(x)) for x in v[:3]))
        lines.append("f -3 -2 -1")
    p = tmp_path / "flake.obj"
    p.write_text("\n".join(lines) + "\n")
    loaded = rt.load_obj(str(p))
    assert np.array_equal(bits(loaded), bits(tris)), "repr(float) round-trips through strtof"
    ob = oracle.OracleBvh(loaded)
    rays = oracle.gen_random_rays(ob.bounds, 4, 0, 50000)
    with rt.Bvh.build(loaded) as gb:
        assert np.array_equal(bits(gb.trace(rays)), bits(ob.trace(rays)))


@pytest.mark.gpu
@pytest.mark.parametrize("key_bits", [30, 63])
def test_bvh_save_load(rt, oracle, tmp_path, key_bits):
    tris = oracle.gen_scene(oracle.SCENE_SOUP, 20000, 0, seed=1)
    ob = oracle.OracleBvh(tris, key_bits)
    rays = oracle.gen_random_rays(ob.bounds, 4, 0, 50000)
    p = str(tmp_path / "scene.bvh")
    with rt.Bvh.build(tris, key_bits=key_bits) as gb:
        gb.save(p)
    with rt.Bvh.load(p) as lb:
        assert lb.n_tris == ob.n and lb.info().key_bits == key_bits
        assert np.array_equal(bits(lb.bounds), bits(ob.bounds))
        left, right, parent = lb.read_topology()
        assert np.array_equal(left, ob.left) and np.array_equal(right, ob.right)
        assert np.array_equal(parent, ob.parent)
        assert np.array_equal(bits(lb.read_boxes()), bits(ob.boxes))
        assert np.array_equal(bits(lb.trace(rays)), bits(ob.trace(rays)))
        lb.widen()
        assert np.array_equal(bits(lb.trace(rays)), bits(ob.trace(rays)))
    for n in (0, 1):
        with rt.Bvh.build(tris[:n]) as gb:
            gb.save(p)
        with rt.Bvh.load(p) as lb:
            assert lb.n_tris == n
            assert np.array_equal(bits(lb.trace(rays[:100])), bits(oracle.OracleBvh(tris[:n]).trace(rays[:100])))
    with open(p, "wb") as f:
        f.write(b"garbage")
    with pytest.raises(rt.RtError):
        rt.Bvh.load(p)
