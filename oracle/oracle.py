"""ctypes front-ends for the two CPU checkers -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

* `PortOracle`  : oracle/raycast_oracle.cpp, the CPU restatement ("port"), built by `build_port()`.
* `RefOracle`   : oracle/_ref/libraycast_ref.so, the reference's own function bodies compiled by
                  oracle/build_ref.py (only buildable where /root/reference exists; the built .so
                  travels to the GPU box).
* `ClRefOracle` : oracle/_ref/libraytracer_cl_ref.so, the reference's OpenCL kernel source compiled as
                  C++ for the host (same recipe); `PortOracle.render(path="opencl")` is its restatement.

Only tests/, __graft_entry__.smoke()/build() and bench.py's cpu_baseline / --impl reference legs
may import this module.  Nothing under opencl_ray_tracer_b200/ does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
PORT_SRC = os.path.join(HERE, "raycast_oracle.cpp")
PORT_SO = os.path.join(HERE, "libraycast_oracle.so")
REF_SO = os.path.join(HERE, "_ref", "libraycast_ref.so")

DEFAULT_O0 = np.array([0, 0, 0, 1], np.float32)
DEFAULT_DU = np.array([1, 0, 0, 0], np.float32)
DEFAULT_DV = np.array([0, 1, 0, 0], np.float32)
DEFAULT_DIR = np.array([0, 0, -1, -1], np.float32)  # states/MainState.cpp:37-39 evaluates to this


def build_port(force=False):
    """g++ -O2 -ffp-contract=off -fopenmp; returns the .so path."""
    if (not force and os.path.isfile(PORT_SO)
            and os.path.getmtime(PORT_SO) >= os.path.getmtime(PORT_SRC)):
        return PORT_SO
    cmd = ["g++", "-std=c++17", "-O2", "-ffp-contract=off", "-fopenmp", "-fPIC", "-shared",
           "-o", PORT_SO, PORT_SRC]
    subprocess.run(cmd, check=True)
    return PORT_SO


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _fp(a):
    return a.ctypes.data_as(C.POINTER(C.c_float)) if a is not None else None


class Scene:
    """Flattened scene arrays, identical to the OpenCL kernel's arguments (rayTracer.cl:111-114)."""

    def __init__(self, sph_o, sph_r, sph_c, cube_v, cube_c):
        self.sph_o = _f32(sph_o).reshape(-1, 4)
        self.sph_r = _f32(sph_r).reshape(-1)
        self.sph_c = _f32(sph_c).reshape(-1, 4)
        self.cube_v = _f32(cube_v).reshape(-1, 36, 4)
        self.cube_c = _f32(cube_c).reshape(-1, 4)
        assert len(self.sph_o) == len(self.sph_r) == len(self.sph_c)
        assert len(self.cube_v) == len(self.cube_c)

    @property
    def ns(self):
        return len(self.sph_r)

    @property
    def nc(self):
        return len(self.cube_c)


class PortOracle:
    def __init__(self):
        self.lib = C.CDLL(build_port())
        L = self.lib
        fp, ip = C.POINTER(C.c_float), C.POINTER(C.c_int32)
        L.orc_render.argtypes = [C.c_int] * 5 + [fp] * 5 + [C.c_int, fp, fp, fp, C.c_int, fp, fp,
                                                            ip, ip, fp, C.c_int, C.c_int]
        L.orc_render.restype = C.c_int
        L.orc_max_threads.restype = C.c_int
        L.orc_fnv1a64.argtypes = [C.c_void_p, C.c_size_t]
        L.orc_fnv1a64.restype = C.c_uint64
        L.orc_fnv1a64_words.argtypes = [C.POINTER(C.c_int32), C.c_size_t]
        L.orc_fnv1a64_words.restype = C.c_uint64
        L.orc_intersect_sphere.argtypes = [fp, fp, C.c_float, fp]
        L.orc_intersect_sphere.restype = C.c_float
        dp = C.POINTER(C.c_double)
        L.orc_intersect_tri.argtypes = [dp] * 6
        L.orc_intersect_tri.restype = C.c_int
        L.orc_pixel_origin.argtypes = [C.c_int, C.c_int, fp, fp, fp, fp]

    def max_threads(self):
        return self.lib.orc_max_threads()

    def render(self, scene, width, height, o0=DEFAULT_O0, du=DEFAULT_DU, dv=DEFAULT_DV,
               ray_dir=DEFAULT_DIR, ray_origins=None, rows=None, faithful=False, threads=0,
               want_t=False, path="cpu"):
        """Returns (rgba_i32[H,W,4], prim_id[H,W], closest[H,W] or None).
        rows = (begin, end, step) renders a subset; untouched rows hold -2 in prim_id.
        path = "cpu" (MainState::executeRayTracerCPU, double triangle test) or "opencl" (the kernel
        rayTracer.cl: float triangle test, OpenCL dot())."""
        n = width * height
        rgba = np.zeros((height, width, 4), np.int32)
        pid = np.full((height, width), -2, np.int32)
        tt = np.zeros((height, width), np.float32) if want_t else None
        o0, du, dv, ray_dir = _f32(o0), _f32(du), _f32(dv), _f32(ray_dir)
        ro = None
        if ray_origins is not None:
            ro = _f32(ray_origins).reshape(n, 4)
        b, e, s = rows if rows is not None else (0, height, 1)
        rc = self.lib.orc_render(
            width, height, b, e, s, _fp(ro), _fp(o0), _fp(du), _fp(dv), _fp(ray_dir),
            scene.ns, _fp(scene.sph_o), _fp(scene.sph_r), _fp(scene.sph_c),
            scene.nc, _fp(scene.cube_v), _fp(scene.cube_c),
            rgba.ctypes.data_as(C.POINTER(C.c_int32)), pid.ctypes.data_as(C.POINTER(C.c_int32)),
            _fp(tt), 2 if path == "opencl" else (1 if faithful else 0), threads)
        if rc != 0:
            raise RuntimeError("orc_render failed: %d" % rc)
        return rgba, pid, tt

    def pixel_origins(self, width, height, o0=DEFAULT_O0, du=DEFAULT_DU, dv=DEFAULT_DV):
        """The float ray origins the oracle uses for the affine camera, shape [H*W, 4]."""
        o0, du, dv = _f32(o0), _f32(du), _f32(dv)
        x = np.arange(width, dtype=np.float32)[None, :, None]
        y = np.arange(height, dtype=np.float32)[:, None, None]
        out = (o0[None, None, :] + x * du[None, None, :]) + y * dv[None, None, :]
        return np.ascontiguousarray(out.astype(np.float32).reshape(-1, 4))

    def fnv1a64(self, arr):
        arr = np.ascontiguousarray(arr)
        return int(self.lib.orc_fnv1a64(arr.ctypes.data, arr.nbytes))


def _fnv_words(self, rgba_i32):
    a = np.ascontiguousarray(rgba_i32, np.int32)
    return int(self.lib.orc_fnv1a64_words(a.ctypes.data_as(C.POINTER(C.c_int32)), a.size))


PortOracle.fnv1a64_words = _fnv_words


def _port_intersect_sphere(self, o4, d4, radius, c4):
    o, d, c = _f32(o4), _f32(d4), _f32(c4)
    return float(self.lib.orc_intersect_sphere(_fp(o), _fp(d), float(radius), _fp(c)))


def _port_intersect_tri(self, o, d, v0, v1, v2):
    arrs = [np.ascontiguousarray(a, np.float64) for a in (o, d, v0, v1, v2)]
    t = np.zeros(1, np.float64)
    dp = C.POINTER(C.c_double)
    hit = self.lib.orc_intersect_tri(*[a.ctypes.data_as(dp) for a in arrs], t.ctypes.data_as(dp))
    return hit, float(t[0])


PortOracle.intersect_sphere = _port_intersect_sphere
PortOracle.intersect_tri = _port_intersect_tri


def narrow_u8(rgba_i32):
    """(uint8_t) narrowing of states/MainState.cpp:1023-1037: value mod 256."""
    return (np.asarray(rgba_i32).astype(np.int64) & 0xFF).astype(np.uint8)


def ref_available():
    return os.path.isfile(REF_SO)


CL_REF_SO = os.path.join(HERE, "_ref", "libraytracer_cl_ref.so")


def cl_ref_available():
    return os.path.isfile(CL_REF_SO)


class ClRefOracle:
    """The reference's OpenCL kernel source (rayTracer.cl) compiled as C++ for the host cores
    (oracle/cl_shim.h, oracle/build_ref.py build_cl): its own arithmetic, one call per work item."""

    def __init__(self):
        if not cl_ref_available():
            raise RuntimeError("oracle/_ref/libraytracer_cl_ref.so missing (python oracle/build_ref.py)")
        self.lib = C.CDLL(CL_REF_SO)
        fp = C.POINTER(C.c_float)
        self.lib.cl_ref_run.argtypes = [C.c_size_t, C.POINTER(C.c_int32), C.c_int, fp, fp, fp, C.c_int, fp, fp, fp, fp,
                                        C.c_int]

    def render(self, scene, origins, ray_dir, threads=0):
        """origins [n,4] (the kernel's rayOrigins argument) -> int32 [n,4] (its `output`)."""
        o = _f32(origins).reshape(-1, 4)
        d = _f32(ray_dir)
        out = np.zeros((len(o), 4), np.int32)
        rc = self.lib.cl_ref_run(len(o), out.ctypes.data_as(C.POINTER(C.c_int32)), scene.ns, _fp(scene.sph_o),
                                 _fp(scene.sph_r), _fp(scene.sph_c), scene.nc, _fp(scene.cube_v), _fp(scene.cube_c),
                                 _fp(o), _fp(d), threads)
        if rc != 0:
            raise RuntimeError("cl_ref_run failed")
        return out


class RefOracle:
    """The reference's own compiled hot path (see oracle/ref_shim.cpp)."""

    def __init__(self):
        if not ref_available():
            raise RuntimeError("oracle/_ref/libraycast_ref.so missing (python oracle/build_ref.py)")
        self.lib = C.CDLL(REF_SO)
        L = self.lib
        fp = C.POINTER(C.c_float)
        L.ref_build_scene.argtypes = [C.c_int, C.c_uint]
        L.ref_scene_counts.argtypes = [C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.ref_scene_export.argtypes = [fp] * 5
        L.ref_add_sphere.argtypes = [fp, C.c_float, fp]
        L.ref_add_cube_ops.argtypes = [fp, C.c_int, C.POINTER(C.c_int), fp, fp]
        L.ref_convert_angle_to_radian.argtypes = [C.c_float]
        L.ref_convert_angle_to_radian.restype = C.c_float
        L.ref_normalise_float.argtypes = [C.c_float] * 3
        L.ref_normalise_float.restype = C.c_float
        L.ref_default_ray_dir.argtypes = [fp]
        L.ref_set_rays.argtypes = [fp, C.c_size_t, fp]
        L.ref_render.argtypes = [C.POINTER(C.c_int32)]
        L.ref_render.restype = C.c_size_t
        L.ref_collide_one.argtypes = [fp, fp, fp]
        L.ref_intersect_sphere.argtypes = [fp, fp, C.c_float, fp]
        L.ref_intersect_sphere.restype = C.c_float
        dp = C.POINTER(C.c_double)
        L.ref_intersect_tri.argtypes = [dp] * 6
        L.ref_intersect_tri.restype = C.c_int
        L.ref_png_decode.argtypes = [C.c_char_p, C.c_void_p, C.c_size_t, C.POINTER(C.c_uint),
                                     C.POINTER(C.c_uint)]
        L.ref_png_encode.argtypes = [C.c_char_p, C.c_void_p, C.c_uint, C.c_uint]

    def export_scene(self):
        ns, nc = C.c_int(), C.c_int()
        self.lib.ref_scene_counts(C.byref(ns), C.byref(nc))
        ns, nc = ns.value, nc.value
        so, sr, sc = (np.zeros((ns, 4), np.float32), np.zeros(ns, np.float32),
                      np.zeros((ns, 4), np.float32))
        cv, cc = np.zeros((nc, 36, 4), np.float32), np.zeros((nc, 4), np.float32)
        self.lib.ref_scene_export(_fp(so), _fp(sr), _fp(sc), _fp(cv), _fp(cc))
        return Scene(so, sr, sc, cv, cc)

    def build_scene(self, which, seed=1):
        if self.lib.ref_build_scene(which, seed) != 0:
            raise ValueError("scene %r" % which)
        return self.export_scene()

    def clear_scene(self):
        self.lib.ref_clear_scene()

    def add_sphere(self, origin4, radius, colour4):
        o, c = _f32(origin4), _f32(colour4)
        self.lib.ref_add_sphere(_fp(o), float(radius), _fp(c))

    def add_cube_ops(self, colour4, ops):
        """ops: list of (kind, (x,y,z)), kind in {'scale','rotate','translate'}; returns verts[36,4]."""
        kinds = np.array([{"scale": 0, "rotate": 1, "translate": 2}[k] for k, _ in ops], np.int32)
        vecs = _f32([v for _, v in ops]).reshape(-1, 3)
        col = _f32(colour4)
        out = np.zeros((36, 4), np.float32)
        self.lib.ref_add_cube_ops(_fp(col), len(ops), kinds.ctypes.data_as(C.POINTER(C.c_int)),
                                  _fp(vecs), _fp(out))
        return out

    def default_ray_dir(self):
        d = np.zeros(4, np.float32)
        self.lib.ref_default_ray_dir(_fp(d))
        return d

    def render(self, origins, ray_dir, shape=None):
        """Runs the reference's ray loop over explicit origins [n,4]; returns int32 [n,4]."""
        o = _f32(origins).reshape(-1, 4)
        d = _f32(ray_dir)
        self.lib.ref_set_rays(_fp(o), len(o), _fp(d))
        out = np.zeros((len(o), 4), np.int32)
        n = self.lib.ref_render(out.ctypes.data_as(C.POINTER(C.c_int32)))
        assert n == out.size
        return out if shape is None else out.reshape(shape)

    def collide_one(self, origin4, dir4):
        o, d, out = _f32(origin4), _f32(dir4), np.zeros(4, np.float32)
        self.lib.ref_collide_one(_fp(o), _fp(d), _fp(out))
        return out

    def intersect_sphere(self, origin4, dir4, radius, centre4):
        o, d, c = _f32(origin4), _f32(dir4), _f32(centre4)
        return float(self.lib.ref_intersect_sphere(_fp(o), _fp(d), float(radius), _fp(c)))

    def intersect_tri(self, o, d, v0, v1, v2):
        arrs = [np.ascontiguousarray(a, np.float64) for a in (o, d, v0, v1, v2)]
        tuv = np.zeros(3, np.float64)
        dp = C.POINTER(C.c_double)
        hit = self.lib.ref_intersect_tri(*[a.ctypes.data_as(dp) for a in arrs],
                                         tuv.ctypes.data_as(dp))
        return hit, tuv

    def png_decode(self, path):
        w, h = C.c_uint(), C.c_uint()
        rc = self.lib.ref_png_decode(path.encode(), None, 0, C.byref(w), C.byref(h))
        if rc != 0:
            raise RuntimeError("lodepng decode error %d" % rc)
        out = np.zeros((h.value, w.value, 4), np.uint8)
        rc = self.lib.ref_png_decode(path.encode(), out.ctypes.data, out.nbytes, C.byref(w), C.byref(h))
        if rc != 0:
            raise RuntimeError("lodepng decode error %d" % rc)
        return out
