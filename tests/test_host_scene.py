"""Host mirror (C++ Cube / Scene / Random / PNG writer) vs the reference's own classes."""
import os
import struct
import zlib

import numpy as np

from parity import GOLDEN


def test_scene1_matches_golden_arrays(built_libs):
    from opencl_ray_tracer_b200 import host_scene as hs
    z = np.load(os.path.join(GOLDEN, "scene1.npz"))
    s = hs.reference_scene(1)
    for name, arr in zip(("sph_o", "sph_r", "sph_c", "cube_v", "cube_c"), s.as_tuple()):
        assert np.array_equal(arr.view(np.uint32), z[name].reshape(arr.shape).view(np.uint32)), name


def test_cube_transforms_bit_exact_vs_reference_cube(ref, built_libs):
    from opencl_ray_tracer_b200 import host_scene as hs
    rng = np.random.default_rng(3)
    for i in range(200):
        ops = [("scale", (rng.uniform(5, 30),) * 3), ("rotate", (0, 0, rng.uniform(0, 6.3))),
               ("rotate", (0, rng.uniform(0, 6.3), 0)), ("rotate", (rng.uniform(0, 6.3), 0, 0)),
               ("translate", tuple(rng.uniform(-100, 8000, 3)))]
        if i % 3 == 0:
            ops.insert(1, ("rotate", tuple(rng.uniform(0, 6.3, 3))))
            ops[0] = ("scale", tuple(rng.uniform(1, 30, 3)))
        ref.clear_scene()
        a = ref.add_cube_ops((1, 0, 0, 255), ops)
        b = hs.cube_from_ops((1, 0, 0, 255), ops)
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
    for a in (0.0, 30.0, 80.0, 250.0, 359.0, 12.345):
        assert ref.lib.ref_convert_angle_to_radian(a) == hs.convert_angle_to_radian(a)


def test_scene2_geometry_matches_reference(ref, built_libs):
    from opencl_ray_tracer_b200 import host_scene as hs
    r = ref.build_scene(2, 1)
    s = hs.reference_scene(2, seed=1)
    assert np.array_equal(r.cube_v.view(np.uint32), s.cube_v.view(np.uint32))  # colours are random, geometry is not
    assert np.array_equal(r.sph_o, s.sph_o) and np.array_equal(r.sph_r, s.sph_r)
    assert np.array_equal(r.cube_c[:4], s.cube_c[:4])


def test_random_scene_recipe_ranges_and_determinism(built_libs):
    from opencl_ray_tracer_b200 import host_scene as hs
    a = hs.random_scene(1920, 1080, 64, 64, seed=2)
    b = hs.random_scene(1920, 1080, 64, 64, seed=2)
    c = hs.random_scene(1920, 1080, 64, 64, seed=3)
    assert all(np.array_equal(x, y) for x, y in zip(a.as_tuple(), b.as_tuple()))
    assert not np.array_equal(a.sph_o, c.sph_o)
    assert a.ns == 64 and a.nc == 64
    assert (a.sph_o[:, 0] >= 0).all() and (a.sph_o[:, 0] <= 1910).all() and (a.sph_o[:, 1] <= 1070).all()
    assert (a.sph_o[:, 2] <= -20).all() and (a.sph_o[:, 2] >= -100).all() and (a.sph_o[:, 3] == 1).all()
    assert (a.sph_r >= 5).all() and (a.sph_r <= 30).all()
    assert (a.sph_c[:, :3] >= 0.05).all() and (a.sph_c[:, :3] <= 1).all() and (a.sph_c[:, 3] == 255).all()
    s3 = hs.reference_scene(3, seed=5)
    assert s3.ns == 100 and s3.nc == 100 and (s3.sph_o[:, 0] <= 630).all()


def _png_decode(path):
    data = open(path, "rb").read()
    assert data[:8] == b"\x89PNG\r\n\x1a\n"
    pos, idat, w, h = 8, b"", 0, 0
    while pos < len(data):
        n, typ = struct.unpack(">I4s", data[pos:pos + 8])
        body = data[pos + 8:pos + 8 + n]
        crc = struct.unpack(">I", data[pos + 8 + n:pos + 12 + n])[0]
        assert zlib.crc32(typ + body) & 0xFFFFFFFF == crc
        if typ == b"IHDR":
            w, h, depth, ctype = struct.unpack(">IIBB", body[:10])
            assert (depth, ctype) == (8, 6)
        if typ == b"IDAT":
            idat += body
        pos += 12 + n
    raw = zlib.decompress(idat)
    rows = np.frombuffer(raw, np.uint8).reshape(h, 1 + 4 * w)
    out = np.zeros((h, 4 * w), np.int32)
    for y in range(h):                      # undo the scanline filters (PNG specification, section 9)
        f, line = int(rows[y, 0]), rows[y, 1:].astype(np.int32)
        up = out[y - 1] if y else np.zeros(4 * w, np.int32)
        cur = out[y]
        assert 0 <= f <= 4
        if f == 0:
            cur[:] = line
        elif f == 2:
            cur[:] = (line + up) & 255
        else:
            for x in range(4 * w):
                a = cur[x - 4] if x >= 4 else 0
                b, c = up[x], (up[x - 4] if x >= 4 else 0)
                if f == 1:
                    pred = a
                elif f == 3:
                    pred = (a + b) >> 1
                else:
                    pp = a + b - c
                    pa, pb, pc = abs(pp - a), abs(pp - b), abs(pp - c)
                    pred = a if (pa <= pb and pa <= pc) else (b if pb <= pc else c)
                cur[x] = (line[x] + pred) & 255
    return out.astype(np.uint8).reshape(h, w, 4)


def test_png_writer_roundtrip(tmp_path, built_libs):
    from opencl_ray_tracer_b200 import host_scene as hs
    img = np.random.default_rng(0).integers(0, 256, (37, 301, 4), dtype=np.uint8)
    p = str(tmp_path / "a.png")
    hs.png_write(p, img)
    assert np.array_equal(_png_decode(p), img)


def test_png_writer_compresses_a_rendered_frame(tmp_path, port, built_libs):
    """Scene 1 at 640x480 (1.2 MB of pixels): real deflate, not stored blocks; decodes to the same pixels."""
    import os
    from opencl_ray_tracer_b200 import host_scene as hs
    from parity import to_oracle_scene
    rgba = port.render(to_oracle_scene(hs.reference_scene(1)), 640, 480)[0]
    img = (rgba & 0xFF).astype(np.uint8)
    p = str(tmp_path / "scene1.png")
    hs.png_write(p, img)
    assert os.path.getsize(p) < img.nbytes // 20
    assert np.array_equal(_png_decode(p), img)
    flat = np.zeros((70, 90, 4), np.uint8)          # long runs: matches of the maximum length, distances of 1 and of a row
    flat[..., 3] = 255
    flat[10:40, 20:60, 0] = 200
    hs.png_write(p, flat)
    assert np.array_equal(_png_decode(p), flat)


def test_png_writer_read_by_reference_lodepng(tmp_path, ref, built_libs):
    from opencl_ray_tracer_b200 import host_scene as hs
    img = np.random.default_rng(1).integers(0, 256, (300, 260, 4), dtype=np.uint8)  # > 65535 B: several stored blocks
    p = str(tmp_path / "b.png")
    hs.png_write(p, img)
    assert np.array_equal(ref.png_decode(p), img)
    grad = np.zeros((200, 333, 4), np.uint8)      # smooth + flat areas: filters 1-4 and LZ77 matches
    grad[..., 0] = np.arange(333)[None, :] % 256
    grad[..., 1] = (np.arange(200)[:, None] * 3) % 256
    grad[50:120, 40:200, 2] = 77
    grad[..., 3] = 255
    hs.png_write(p, grad)
    assert np.array_equal(ref.png_decode(p), grad)
