"""Parity metric + helpers shared by the CPU and GPU tests.

ID image : identical on every pixel that is not a silhouette pixel (a pixel whose ORACLE id
           differs from any of its 8 neighbours' oracle ids).
RGB      : |delta| <= 1 per channel (mod-256 values, i.e. after the reference's uint8 narrowing,
           compared cyclically so 255 vs 0 across a wrap counts as 1) on >= 99.9 % of pixels;
           every remaining pixel must be a silhouette pixel.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
GOLDEN = os.path.join(HERE, "golden")


def silhouette_mask(ids):
    h, w = ids.shape
    pad = np.pad(ids, 1, mode="edge")
    sil = np.zeros((h, w), bool)
    for dy in range(3):
        for dx in range(3):
            sil |= pad[dy:dy + h, dx:dx + w] != ids
    return sil


def narrow(rgba_i32):
    return (np.asarray(rgba_i32).astype(np.int64) & 0xFF).astype(np.int32)


def parity_report(oracle_rgba_i32, oracle_ids, got_rgba_i32, got_ids):
    sil = silhouette_mask(oracle_ids)
    id_bad = oracle_ids != got_ids
    a, b = narrow(oracle_rgba_i32)[..., :3], narrow(got_rgba_i32)[..., :3]
    d = np.abs(a - b)
    d = np.minimum(d, 256 - d).max(-1)  # cyclic distance of the wrapped bytes
    n = oracle_ids.size
    return {
        "pixels": n,
        "silhouette_pixels": int(sil.sum()),
        "id_mismatch": int(id_bad.sum()),
        "id_mismatch_off_silhouette": int((id_bad & ~sil).sum()),
        "rgb_exact_frac": float((d == 0).mean()),
        "rgb_within1_frac": float((d <= 1).mean()),
        "rgb_gt1": int((d > 1).sum()),
        "rgb_gt1_off_silhouette": int(((d > 1) & ~sil).sum()),
        "alpha_bad": int((narrow(got_rgba_i32)[..., 3] != 255).sum()),
    }


def assert_parity(rep, min_within1=0.999):
    assert rep["id_mismatch_off_silhouette"] == 0, rep
    assert rep["rgb_gt1_off_silhouette"] == 0, rep
    assert rep["rgb_within1_frac"] >= min_within1, rep
    assert rep["alpha_bad"] == 0, rep


def default_origins(w, h, x0=0, y0=0):
    xs, ys = np.meshgrid(np.arange(x0, x0 + w, dtype=np.float32), np.arange(y0, y0 + h, dtype=np.float32))
    o = np.zeros((h, w, 4), np.float32)
    o[..., 0], o[..., 1], o[..., 3] = xs, ys, 1.0
    return o.reshape(-1, 4)


def to_oracle_scene(s):
    from oracle.oracle import Scene
    return Scene(s.sph_o, s.sph_r, s.sph_c, s.cube_v, s.cube_c)


def load_npz_scene(z, prefix=""):
    from oracle.oracle import Scene
    return Scene(z[prefix + "sph_o"], z[prefix + "sph_r"], z[prefix + "sph_c"], z[prefix + "cube_v"],
                 z[prefix + "cube_c"])


def kat_names():
    z = np.load(os.path.join(GOLDEN, "kats.npz"))
    return sorted({k.split("/")[0] for k in z.files})


class Emulator:
    """tests/emu/emu_raycast.cpp: scalar CPU walk-through of the kernel arithmetic (test-only)."""

    def __init__(self):
        src = os.path.join(HERE, "emu", "emu_raycast.cpp")
        so = os.path.join(HERE, "emu", "libemu_raycast.so")
        hdr = os.path.join(ROOT, "opencl_ray_tracer_b200", "csrc", "raycast_math.h")
        if not os.path.isfile(so) or os.path.getmtime(so) < max(os.path.getmtime(src), os.path.getmtime(hdr)):
            subprocess.run(["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-fopenmp", "-fPIC", "-shared",
                            "-o", so, src], check=True)
        self.lib = C.CDLL(so)

    def render(self, scene, w, h, camera=None):
        from opencl_ray_tracer_b200 import capi
        camera = camera or capi.REFERENCE_CAMERA
        cr, sr, cam = capi.debug_pack(scene.sph_o, scene.sph_r, scene.cube_v, camera)
        rgba = np.zeros((h, w, 4), np.int32)
        pid = np.zeros((h, w), np.int32)
        tt = np.zeros((h, w), np.float32)
        fp = lambda a: np.ascontiguousarray(a, np.float32).ctypes.data_as(C.POINTER(C.c_float))
        ip = lambda a: a.ctypes.data_as(C.POINTER(C.c_int32))
        cc, sc = np.ascontiguousarray(scene.cube_c, np.float32), np.ascontiguousarray(scene.sph_c, np.float32)
        self.lib.emu_render_grid(w, h, fp(cam), scene.nc, fp(cr), scene.ns, fp(sr), fp(cc), fp(sc),
                                 ip(rgba), ip(pid), fp(tt))
        return rgba, pid, tt

    def check_tile_cull(self, scene, w, h, camera=None, tile_w=256, tile_h=8):
        """Walks raycast_cull_kernel's tile test: dict(violations, culled, empty, total) over (tile, primitive) pairs."""
        from opencl_ray_tracer_b200 import capi
        camera = camera or capi.REFERENCE_CAMERA
        _, _, cam = capi.debug_pack(scene.sph_o, scene.sph_r, scene.cube_v, camera)
        bnd = np.ascontiguousarray(capi.debug_pack_bounds(scene.sph_o, scene.sph_r, scene.cube_v, w, h, camera), np.float32)
        fp = lambda a: np.ascontiguousarray(a, np.float32).ctypes.data_as(C.POINTER(C.c_float))
        out = (C.c_longlong * 4)()
        self.lib.emu_check_tile_cull(w, h, fp(cam), len(bnd), fp(bnd), tile_w, tile_h, out)
        return {"violations": out[0], "culled": out[1], "empty": out[2], "total": out[3]}

    def check_div180(self, first_bits, last_bits, stride=1):
        """rtb_div180 vs the IEEE divide on every stride-th float bit pattern in [first_bits, last_bits]: mismatches."""
        self.lib.emu_check_div180.restype = C.c_longlong
        self.lib.emu_check_div180.argtypes = [C.c_uint, C.c_uint, C.c_uint]
        return int(self.lib.emu_check_div180(first_bits, last_bits, stride))

    def check_min_form(self, scene, w, h, camera=None, run=16):
        """fmaf(min|w|, min|w|, rowC) vs the minimum of the per-ray q over runs of `run` pixels: number of differences."""
        from opencl_ray_tracer_b200 import capi
        camera = camera or capi.REFERENCE_CAMERA
        _, _, cam = capi.debug_pack(scene.sph_o, scene.sph_r, scene.cube_v, camera)
        bnd = np.ascontiguousarray(capi.debug_pack_bounds(scene.sph_o, scene.sph_r, scene.cube_v, w, h, camera), np.float32)
        fp = lambda a: np.ascontiguousarray(a, np.float32).ctypes.data_as(C.POINTER(C.c_float))
        self.lib.emu_check_min_form.restype = C.c_longlong
        return int(self.lib.emu_check_min_form(w, h, fp(cam), len(bnd), fp(bnd), run))

    def check_row_table(self, scene, w, h, camera=None):
        """The dense kernel's row-table test against the per-ray q test: dict(violations, pass_q, pass_row, total)."""
        from opencl_ray_tracer_b200 import capi
        camera = camera or capi.REFERENCE_CAMERA
        _, _, cam = capi.debug_pack(scene.sph_o, scene.sph_r, scene.cube_v, camera)
        bnd = np.ascontiguousarray(capi.debug_pack_bounds(scene.sph_o, scene.sph_r, scene.cube_v, w, h, camera), np.float32)
        tab = np.ascontiguousarray(capi.debug_pack_row_table(bnd, w, h, camera), np.float32)
        fp = lambda a: np.ascontiguousarray(a, np.float32).ctypes.data_as(C.POINTER(C.c_float))
        out = (C.c_longlong * 4)()
        self.lib.emu_check_row_table(w, h, fp(cam), len(bnd), fp(bnd), fp(tab), out)
        return {"violations": out[0], "pass_q": out[1], "pass_row": out[2], "total": out[3]}

    def check_bounds(self, scene, w, h, camera=None):
        """Walks the kernel's dense loop: returns dict(violations, bound_pass, footprint, total)."""
        from opencl_ray_tracer_b200 import capi
        camera = camera or capi.REFERENCE_CAMERA
        cr, sr, cam = capi.debug_pack(scene.sph_o, scene.sph_r, scene.cube_v, camera)
        bnd = capi.debug_pack_bounds(scene.sph_o, scene.sph_r, scene.cube_v, w, h, camera)
        fp = lambda a: np.ascontiguousarray(a, np.float32).ctypes.data_as(C.POINTER(C.c_float))
        out = (C.c_longlong * 4)()
        self.lib.emu_check_bounds(w, h, fp(cam), scene.nc, fp(cr), scene.ns, fp(sr), fp(bnd), out)
        return {"violations": out[0], "bound_pass": out[1], "footprint": out[2], "total": out[3]}
