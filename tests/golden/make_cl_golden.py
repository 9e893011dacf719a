#!/usr/bin/env python3
"""Golden vectors of the reference's OTHER path, the OpenCL kernel `rayTracer`
(RayTrace/resources/shaders/rayTracer.cl), produced by the kernel's OWN source compiled as C++ for the
host (oracle/_ref/libraytracer_cl_ref.so: oracle/build_ref.py build_cl + oracle/cl_shim.h; no OpenCL
ICD exists in this image).  Run in the build container (needs /root/reference):
    python tests/golden/make_cl_golden.py
Outputs (small, committed): tests/golden/cl_path.npz, tests/golden/cl_path.json.  The scenes are the
ones of the CPU-path golden files (scene1.npz, the scene-3 window, kats.npz), so the two paths can be
compared pixel by pixel.
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), "..", ".."))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle import build_ref  # noqa: E402
from oracle.oracle import DEFAULT_DIR, ClRefOracle, PortOracle  # noqa: E402
from parity import default_origins, kat_names, load_npz_scene  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))


def main():
    build_ref.build_cl()
    cl, port = ClRefOracle(), PortOracle()
    out, info = {}, {}
    sc1 = load_npz_scene(np.load(os.path.join(OUT, "scene1.npz")))
    full = cl.render(sc1, default_origins(640, 480), DEFAULT_DIR).reshape(480, 640, 4)
    cpu = port.render(sc1, 640, 480)[0]
    info["scene1_640x480"] = {
        "fnv1a64_words": "%016x" % port.fnv1a64_words(full),
        "pixels_with_nonzero_rgb": int((full[..., :3] != 0).any(-1).sum()),
        "sum_of_all_ints": int(full.astype(np.int64).sum()),
        "pixels_differing_from_cpu_path": int((full != cpu).any(-1).sum()),
    }
    out["scene1/sub4"] = full[::4, ::4].copy()
    ys, xs = np.nonzero((full != cpu).any(-1))          # every pixel on which the two reference paths disagree
    out["scene1/diff_yx"] = np.stack([ys, xs], 1).astype(np.int32)
    out["scene1/diff_rgba"] = full[ys, xs].copy()
    z3 = np.load(os.path.join(OUT, "scene3_seed7_window.npz"))
    sc3 = load_npz_scene(z3)
    h, w = z3["rgba"].shape[:2]
    out["scene3_seed7_window/rgba"] = cl.render(sc3, default_origins(w, h, int(z3["x0"]), int(z3["y0"])), DEFAULT_DIR).reshape(h, w, 4)
    zk = np.load(os.path.join(OUT, "kats.npz"))
    for name in kat_names():
        sc = load_npz_scene(zk, name + "/")
        out["kat/" + name] = cl.render(sc, default_origins(64, 48), DEFAULT_DIR).reshape(48, 64, 4)
    np.savez_compressed(os.path.join(OUT, "cl_path.npz"), **out)
    json.dump(info, open(os.path.join(OUT, "cl_path.json"), "w"), indent=1)
    print(json.dumps(info))


if __name__ == "__main__":
    main()
