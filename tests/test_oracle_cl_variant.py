"""The reference's OTHER path -- the OpenCL kernel `rayTracer` (rayTracer.cl:111-202: float triangle
test, OpenCL dot()) -- as a second parity target.

* the port's `path="opencl"` restatement is held bit for bit to the committed golden vectors made from
  the kernel's own source compiled as C++ (tests/golden/make_cl_golden.py), and, where oracle/_ref
  exists, to that compiled kernel live on further scenes;
* the two reference paths themselves differ only on silhouette pixels (33 of scene 1's 307 200).
"""
import json
import os

import numpy as np
import pytest

from parity import GOLDEN, assert_parity, default_origins, kat_names, load_npz_scene, parity_report, to_oracle_scene

DIR = np.array([0, 0, -1, -1], np.float32)


def test_scene1_known_answer_of_the_opencl_path(port):
    info = json.load(open(os.path.join(GOLDEN, "cl_path.json")))["scene1_640x480"]
    z = np.load(os.path.join(GOLDEN, "cl_path.npz"))
    sc = load_npz_scene(np.load(os.path.join(GOLDEN, "scene1.npz")))
    rgba, pid, _ = port.render(sc, 640, 480, path="opencl")
    assert "%016x" % port.fnv1a64_words(rgba) == info["fnv1a64_words"] == "6cc2d72c7f0b365f"
    assert int((rgba[..., :3] != 0).any(-1).sum()) == info["pixels_with_nonzero_rgb"] == 38311
    assert int(rgba.astype(np.int64).sum()) == info["sum_of_all_ints"]
    assert np.array_equal(rgba[::4, ::4], z["scene1/sub4"])
    # the pixels on which the reference's two paths disagree, and what the OpenCL path has there
    cpu, cpu_id, _ = port.render(sc, 640, 480)
    ys, xs = np.nonzero((rgba != cpu).any(-1))
    assert np.array_equal(np.stack([ys, xs], 1), z["scene1/diff_yx"]) and len(ys) == info["pixels_differing_from_cpu_path"] == 33
    assert np.array_equal(rgba[ys, xs], z["scene1/diff_rgba"])
    rep = parity_report(cpu, cpu_id, rgba, pid)      # ... all of them silhouette pixels
    assert_parity(rep)
    assert rep["id_mismatch"] == 29


def test_scene3_window_and_kats_of_the_opencl_path(port):
    z = np.load(os.path.join(GOLDEN, "cl_path.npz"))
    z3 = np.load(os.path.join(GOLDEN, "scene3_seed7_window.npz"))
    sc3 = load_npz_scene(z3)
    h, w = z3["rgba"].shape[:2]
    got = port.render(sc3, w, h, ray_origins=default_origins(w, h, int(z3["x0"]), int(z3["y0"])), ray_dir=DIR, path="opencl")[0]
    assert np.array_equal(got, z["scene3_seed7_window/rgba"])
    zk = np.load(os.path.join(GOLDEN, "kats.npz"))
    for name in kat_names():
        sc = load_npz_scene(zk, name + "/")
        assert np.array_equal(port.render(sc, 64, 48, path="opencl")[0], z["kat/" + name]), name


@pytest.mark.parametrize("w,h,ns,nc,seed", [(320, 240, 60, 60, 5), (256, 128, 0, 40, 6), (200, 150, 90, 0, 8)])
def test_port_equals_compiled_kernel_on_random_scenes(port, cl_ref, built_libs, w, h, ns, nc, seed):
    from opencl_ray_tracer_b200 import host_scene as hs
    sc = to_oracle_scene(hs.random_scene(w, h, ns, nc, seed))
    a = port.render(sc, w, h, path="opencl")[0]
    b = cl_ref.render(sc, default_origins(w, h), DIR).reshape(h, w, 4)
    assert np.array_equal(a, b)
