"""CPU tests of host logic added in round 2: the shard rule in C and in Python, the image
writer against a committed golden frame (§8f.1), the C99 view of the header, and the
oracle's box test for rays with zero direction components against brute force."""
import hashlib
import json
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import bits

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")


def test_shard_range_c_equals_python(rt):
    from gpu_raytracer_b200 import sharding
    for n in (0, 1, 31, 32, 33, 1000, 1_000_003, 10**9):
        for world in (1, 2, 3, 4, 8):
            got = [rt.shard_range(n, r, world) for r in range(world)]
            assert got == [sharding.shard_range(n, r, world) for r in range(world)]
            assert got[0][0] == 0 and sum(c for _, c in got) == n
            for (f0, c0), (f1, _) in zip(got, got[1:]):
                assert f0 + c0 == f1
    with pytest.raises(rt.RtError):
        rt.shard_range(10, 3, 3)


def test_quantize_is_the_gamma_curve(rt):
    q = rt.image.quantize
    assert q(np.float32(0.0)) == 0 and q(np.float32(1.0)) == 255 and q(np.float32(2.0)) == 255
    assert q(np.float32(-1.0)) == 0 and q(np.float32("nan")) == 0
    x = np.linspace(0, 1, 4001, dtype=np.float32)
    want = np.floor(x.astype(np.float64) ** (1 / 2.2) * 255.0 + 0.5)
    assert np.abs(q(x).astype(np.int64) - want.astype(np.int64)).max() == 0
    assert (np.diff(q(x).astype(np.int64)) >= 0).all()


def test_write_ppm_golden_frame(rt, oracle, tmp_path):
    """The Cornell frame of tests/golden/make_golden.py: oracle trace + AO + shade, written
    with image.write_ppm; header, byte count and sha256 are pinned."""
    sys.path.insert(0, GOLDEN)
    import make_golden
    want = json.load(open(os.path.join(GOLDEN, "cornell_96_frame.json")))
    rgb, radius = make_golden.cornell_frame(want["width"], want["height"], want["spp"], want["seed"])
    assert radius == want["radius"]
    assert hashlib.sha256(rgb.tobytes()).hexdigest() == want["rgb_sha256"]
    path = tmp_path / "frame.ppm"
    rt.image.write_ppm(str(path), rgb, want["width"], want["height"])
    data = path.read_bytes()
    header = f"P6\n{want['width']} {want['height']}\n255\n".encode()
    assert data.startswith(header)
    assert len(data) == len(header) + 3 * want["width"] * want["height"] == want["ppm_bytes"]
    assert hashlib.sha256(data).hexdigest() == want["ppm_sha256"]
    px = np.frombuffer(data[len(header):], dtype=np.uint8).reshape(want["height"], want["width"], 3)
    assert (px[..., 0] == px[..., 1]).all() and px.max() > 200 and (px == 0).any()


def test_write_png_holds_the_ppm_pixels(rt, tmp_path):
    """image.write_png: a valid 8-bit RGB PNG (signature, IHDR, CRCs) whose decoded pixels are
    write_ppm's; checked with the standard library, and with Pillow where it is installed."""
    import struct
    import zlib
    rng = np.random.default_rng(3)
    w, h = 37, 21
    rgb = rng.random((w * h, 3)).astype(np.float32) * 1.2 - 0.1
    rt.image.write_ppm(str(tmp_path / "a.ppm"), rgb, w, h)
    rt.image.write_png(str(tmp_path / "a.png"), rgb, w, h)
    ppm = (tmp_path / "a.ppm").read_bytes()
    want = np.frombuffer(ppm[len(f"P6\n{w} {h}\n255\n"):], dtype=np.uint8).reshape(h, w, 3)
    data = (tmp_path / "a.png").read_bytes()
    assert data[:8] == b"\x89PNG\r\n\x1a\n"
    at, chunks = 8, []
    while at < len(data):
        (n,) = struct.unpack(">I", data[at:at + 4])
        tag, body = data[at + 4:at + 8], data[at + 8:at + 8 + n]
        assert struct.unpack(">I", data[at + 8 + n:at + 12 + n])[0] == zlib.crc32(tag + body) & 0xFFFFFFFF
        chunks.append((tag, body))
        at += 12 + n
    assert [t for t, _ in chunks] == [b"IHDR", b"IDAT", b"IEND"]
    assert struct.unpack(">IIBBBBB", chunks[0][1]) == (w, h, 8, 2, 0, 0, 0)
    rows = np.frombuffer(zlib.decompress(chunks[1][1]), dtype=np.uint8).reshape(h, 1 + 3 * w)
    assert (rows[:, 0] == 0).all() and np.array_equal(rows[:, 1:].reshape(h, w, 3), want)
    try:
        from PIL import Image
    except ImportError:
        return
    assert np.array_equal(np.asarray(Image.open(str(tmp_path / "a.png")).convert("RGB")), want)


def test_header_compiles_and_links_as_c99(rt, tmp_path):
    """include/gpu_raytracer.h is a C header: a C99 translation unit includes it, links against
    the library and drives build_bvh / trace_rays / free_bvh (tests/c99/smoke.c)."""
    exe = tmp_path / "smoke"
    lib_dir = os.path.dirname(rt.SO_PATH)
    cmd = ["gcc", "-std=c99", "-Wall", "-Wextra", "-Werror", "-pedantic", "-I", os.path.join(ROOT, "include"),
           os.path.join(ROOT, "tests", "c99", "smoke.c"), "-o", str(exe), rt.SO_PATH,
           f"-Wl,-rpath,{lib_dir}", "-lm"]
    subprocess.run(cmd, check=True, env={**os.environ, "CC": "gcc"})
    res = subprocess.run([str(exe)], capture_output=True, text=True)
    assert res.returncode == 0, res.stdout + res.stderr
    assert ("no device" in res.stdout) or ("ok" in res.stdout)


def test_cpp_bench_main_compiles_and_links(rt, tmp_path):
    """tools/cpp/bench_main.cpp (SURVEY.md §8b "a small C++ bench main") builds against the header
    and the library with nothing from Python in it; without a device it says so and exits 0."""
    import json
    exe = tmp_path / "bench_main"
    lib_dir = os.path.dirname(rt.SO_PATH)
    cuda = os.environ.get("CUDA_HOME", "/usr/local/cuda")
    subprocess.run(["g++", "-std=c++17", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), "-I", os.path.join(cuda, "include"),
                    os.path.join(ROOT, "tools", "cpp", "bench_main.cpp"), "-o", str(exe), rt.SO_PATH,
                    "-L", os.path.join(cuda, "lib64"), "-lcudart", f"-Wl,-rpath,{lib_dir}", f"-Wl,-rpath,{os.path.join(cuda, 'lib64')}"],
                   check=True)
    res = subprocess.run([str(exe), "cornell", "64", "64", "1"], capture_output=True, text=True)
    assert res.returncode == 0, res.stdout + res.stderr
    out = json.loads(res.stdout)
    assert out.get("error") == "no CUDA device" or out["n_tris"] == 1024


def _grid_plane(k=8):
    """[0,1]^2 at z = 0 as k x k x 2 triangles on exact binary fractions."""
    t = []
    s = 1.0 / k
    for i in range(k):
        for j in range(k):
            x0, y0, x1, y1 = i * s, j * s, (i + 1) * s, (j + 1) * s
            t.append([x0, y0, 0, 0, x1, y0, 0, 0, x1, y1, 0, 0])
            t.append([x0, y0, 0, 0, x1, y1, 0, 0, x0, y1, 0, 0])
    return np.array(t, dtype=np.float32)


def zero_direction_cases():
    """(tris, rays): rays with +-0 direction components whose origins lie on box faces, grid
    lines and vertices: the cases a clamped 1/d got wrong."""
    tris = np.concatenate([_grid_plane(8), _grid_plane(8) + np.float32([0, 0, 2, 0] * 3)])
    rays = []
    for pz, nz in ((0.0, 1.0), (-0.0, 1.0), (0.0, -1.0)):
        for x in (0.0, 0.125, 0.5, 0.3, 1.0, 0.0625):
            for y in (0.0, 0.25, 0.5, 0.7, 1.0, 0.0625):
                for oz in (-1.0, 1.0, 3.0):
                    rays.append([x, y, oz, 0, pz, -pz, nz, np.inf])
                    rays.append([x, y, oz, 0, -pz, pz, nz, np.inf])
    # rays lying IN the planes z = 0 and z = 2, and along box edges
    for y in (0.0, 0.5, 1.0):
        rays.append([-1.0, y, 0.0, 0, 1.0, 0.0, 0.0, np.inf])
        rays.append([-1.0, y, 2.0, 0, 1.0, -0.0, 0.0, np.inf])
        rays.append([0.25, -1.0, 1.0, 0, 0.0, 1.0, 0.0, np.inf])
    # the advisor's case: origin on the hi face of a child box, d = (+0, 0, 1)
    rays.append([0.5, 0.25, -1.0, 0, 0.0, 0.0, 1.0, np.inf])
    rays.append([0.5, 0.25, -1.0, 0, -0.0, 0.0, 1.0, np.inf])
    return tris, np.array(rays, dtype=np.float32)


def test_oracle_zero_direction_rays_match_brute_force(oracle):
    tris, rays = zero_direction_cases()
    for key_bits in (30, 63):
        b = oracle.OracleBvh(tris, key_bits)
        for flags in (oracle.TRACE_WATERTIGHT, oracle.TRACE_CLOSEST, oracle.TRACE_ANY | oracle.TRACE_WATERTIGHT):
            got, want = b.trace(rays, flags), oracle.brute(tris, rays, flags)
            assert np.array_equal(bits(got), bits(want)), (key_bits, flags)
    hit = oracle.hit_prims(oracle.brute(tris, rays, oracle.TRACE_WATERTIGHT)) >= 0
    assert hit.sum() > 100 and (~hit).sum() > 3
    # the advisor's scene (2 triangles, one spanning x in [0, 0.5])
    t = np.zeros((2, 12), dtype=np.float32)
    t[0] = [0, 0, 0, 0, 0.5, 0, 0, 0, 0.5, 0.5, 0, 0]
    t[1] = [2, 2, 2, 0, 3, 2, 2, 0, 3, 3, 2, 0]
    r = np.array([[0.5, 0.25, -1, 0, 0.0, 0, 1, np.inf], [0.5, 0.25, -1, 0, -0.0, 0, 1, np.inf]], dtype=np.float32)
    got = oracle.OracleBvh(t).trace(r)
    assert oracle.hit_prims(got).tolist() == [0, 0] and got[:, 0].tolist() == [1.0, 1.0]


def test_oracle_shade(oracle):
    tri = np.array([[0, 0, 1, 0, 1, 0, 1, 0, 0, 1, 1, 0]], dtype=np.float32)
    rays = np.array([[0.25, 0.25, 0, 0, 0, 0, 1, np.inf], [5, 5, 0, 0, 0, 0, 1, np.inf]], dtype=np.float32)
    hits = oracle.OracleBvh(tri).trace(rays)
    rgb = oracle.shade(tri, rays, hits)
    assert rgb.tolist() == [[1.0, 1.0, 1.0], [0.0, 0.0, 0.0]]
    occ = np.zeros((4, 4), dtype=np.float32)
    occ[:, 3] = np.array([0x7FFFFFFF, -1, -1, -1], dtype=np.int32).view(np.float32)
    assert oracle.shade(tri, rays, hits, occ, 2)[0, 0] == 0.5
