#!/usr/bin/env python3
"""Generates the committed golden vectors from oracle/_ref = the REFERENCE'S OWN compiled functions
(MainState::collide / intersectTri / intersectSphere / createScene1-3 / the executeRayTracerCPU
loop, RayTrace/Cube.cpp, glm) -- see oracle/build_ref.py.  Run in the build container (needs
/root/reference):   python tests/golden/make_golden.py
Outputs (small, committed): tests/golden/*.npz, tests/golden/scene1_640x480.json
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), "..", ".."))
sys.path.insert(0, ROOT)
from oracle import build_ref  # noqa: E402
from oracle.oracle import PortOracle, RefOracle, Scene  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def grid(w, h, x0=0, y0=0):
    xs, ys = np.meshgrid(np.arange(x0, x0 + w, dtype=np.float32), np.arange(y0, y0 + h, dtype=np.float32))
    o = np.zeros((h, w, 4), np.float32)
    o[..., 0], o[..., 1], o[..., 3] = xs, ys, 1.0
    return o.reshape(-1, 4)


def scene_dict(sc):
    return dict(sph_o=sc.sph_o, sph_r=sc.sph_r, sph_c=sc.sph_c, cube_v=sc.cube_v, cube_c=sc.cube_c)


def main():
    build_ref.build()
    ref, port = RefOracle(), PortOracle()
    d = ref.default_ray_dir()
    assert d.tolist() == [0.0, 0.0, -1.0, -1.0]

    # 1. scene 1, the reference's default 640x480 frame: statistics + hash of the int stream
    sc1 = ref.build_scene(1)
    full = ref.render(grid(640, 480), d, (480, 640, 4))
    lit = int(((full[..., :3] != 0).any(-1)).sum())
    info = {
        "width": 640, "height": 480, "ray_dir": d.tolist(),
        "fnv1a64_words": "%016x" % port.fnv1a64_words(full),
        "fnv1a64_bytes": "%016x" % port.fnv1a64(full),
        "pixels_with_nonzero_rgb": lit,
        "pixels_with_channel_gt_255": int((full[..., :3] > 255).any(-1).sum()),
        "sum_of_all_ints": int(full.astype(np.int64).sum()),
    }
    json.dump(info, open(os.path.join(OUT, "scene1_640x480.json"), "w"), indent=1)
    # every 4th row/column of that frame, exact ints, plus the scene itself
    np.savez_compressed(os.path.join(OUT, "scene1.npz"), sub4=full[::4, ::4].copy(), **scene_dict(sc1))

    # 2. scene 3 (100 spheres + 100 cubes, libc rand seed 7): a 160x96 window at (200,150)
    sc3 = ref.build_scene(3, 7)
    win = ref.render(grid(160, 96, 200, 150), d, (96, 160, 4))
    np.savez_compressed(os.path.join(OUT, "scene3_seed7_window.npz"), x0=200, y0=150, rgba=win, **scene_dict(sc3))

    # 3. scene 2 (libc rand seed 3): 128x96 window at (380,120)
    sc2 = ref.build_scene(2, 3)
    win2 = ref.render(grid(128, 96, 380, 120), d, (96, 128, 4))
    np.savez_compressed(os.path.join(OUT, "scene2_seed3_window.npz"), x0=380, y0=120, rgba=win2, **scene_dict(sc2))

    # 4. known-answer cases, each a tiny scene rendered on a 64x48 grid by the reference code
    kats = {}

    def kat(name, spheres=(), cubes=()):
        ref.clear_scene()
        for o, r, c in spheres:
            ref.add_sphere(o, r, c)
        for col, ops in cubes:
            ref.add_cube_ops(col, ops)
        sc = ref.export_scene()
        img = ref.render(grid(64, 48), d, (48, 64, 4))
        kats[name] = dict(rgba=img, **scene_dict(sc))

    R, G, B, Y = (1, 0, 0, 255), (0, 1, 0, 255), (0, 0, 1, 255), (1, 1, 0, 255)
    kat("axis_aligned_cube_pixel_exact_edges", cubes=[(R, [("scale", (10, 10, 10)), ("translate", (30, 20, -70))])])
    kat("sphere_origin_inside_negative_t0", spheres=[((32, 24, -5, 1), 20, G)])
    kat("sphere_behind_plane_tca_negative", spheres=[((32, 24, 30, 1), 10, G)])
    kat("sphere_t0_exactly_zero_is_a_miss", spheres=[((32, 24, -10, 1), 10, G), ((32, 24, -50, 1), 12, B)])
    kat("cube_behind_plane_negative_t_wins", cubes=[(R, [("scale", (8, 8, 8)), ("translate", (20, 20, 40))])],
        spheres=[((24, 22, -60, 1), 15, G)])
    kat("equal_t_cube_before_sphere", cubes=[(R, [("scale", (10, 10, 10)), ("translate", (30, 24, -60))])],
        spheres=[((36, 24, -60, 1), 10, G)])
    kat("equal_t_lower_index_wins", spheres=[((30, 24, -60, 1), 12, G), ((30, 24, -60, 1), 12, B)],
        cubes=[(R, [("scale", (6, 6, 6)), ("translate", (50, 10, -40))]), (Y, [("scale", (6, 6, 6)), ("translate", (50, 10, -40))])])
    kat("t_beyond_180_wraps", spheres=[((20, 24, -260, 1), 18, G)],
        cubes=[(B, [("scale", (9, 9, 9)), ("rotate", (0.3, 0.5, 0.7)), ("translate", (45, 24, -300))])])
    kat("rotated_cube_three_rotations", cubes=[(Y, [("scale", (14, 14, 14)), ("rotate", (0, 0, 0.6)), ("rotate", (0, 1.1, 0)),
                                                    ("rotate", (2.0, 0, 0)), ("translate", (32, 24, -50))])])
    kat("sheared_parallelepiped", cubes=[(Y, [("rotate", (0.4, 0.2, 0.9)), ("scale", (16, 7, 11)), ("translate", (30, 22, -50))])])
    kat("empty_scene")
    flat = {}
    for k, v in kats.items():
        for kk, vv in v.items():
            flat[k + "/" + kk] = vv
    np.savez_compressed(os.path.join(OUT, "kats.npz"), **flat)

    # 5. single-function vectors: intersectSphere / intersectTri / collide on hand-picked inputs
    rng = np.random.default_rng(12345)
    sph_in = np.zeros((64, 13), np.float32)
    sph_out = np.zeros(64, np.float32)
    for i in range(64):
        o = np.array([rng.uniform(0, 640), rng.uniform(0, 480), 0, 1], np.float32)
        c = np.array([o[0] + rng.uniform(-30, 30), o[1] + rng.uniform(-30, 30), rng.uniform(-100, 20), 1], np.float32)
        r = np.float32(rng.uniform(5, 30))
        sph_in[i] = np.concatenate([o, d, [r], c])
        sph_out[i] = ref.intersect_sphere(o, d, r, c)
    tri_in = np.zeros((64, 15), np.float64)
    tri_out = np.zeros((64, 4), np.float64)
    for i in range(64):
        o = np.array([rng.uniform(0, 64), rng.uniform(0, 48), 0.0])
        dd = np.array([0.0, 0.0, -1.0])
        v = rng.uniform(-20, 80, (3, 3)).astype(np.float32).astype(np.float64)
        v[:, 2] = rng.uniform(-90, 10, 3)
        hit, tuv = ref.intersect_tri(o, dd, v[0], v[1], v[2])
        tri_in[i] = np.concatenate([o, dd, v.reshape(-1)])
        tri_out[i] = [hit, tuv[0] if hit else 0, tuv[1] if hit else 0, tuv[2] if hit else 0]
    np.savez_compressed(os.path.join(OUT, "function_vectors.npz"), sph_in=sph_in, sph_out=sph_out,
                        tri_in=tri_in, tri_out=tri_out)
    print("golden vectors written to", OUT)
    print(json.dumps(info))


if __name__ == "__main__":
    main()
