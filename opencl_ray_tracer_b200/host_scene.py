"""ctypes binding of librtb_host.so: the C++ host mirror's scene builders (createScene1/2/3 and the
seeded synthetic generator), orbit camera and PNG writer, so Python drivers use exactly the scenes
the C++ driver (raycast_cli) uses.  Pure host code."""
import ctypes as C
import os

import numpy as np

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
HOST_LIB_PATH = os.path.join(PKG_DIR, "librtb_host.so")

_lib = None


def load_host_library():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.isfile(HOST_LIB_PATH):
        raise RuntimeError("librtb_host.so not found -- run opencl_ray_tracer_b200/build.sh")
    lib = C.CDLL(HOST_LIB_PATH)
    fp = C.POINTER(C.c_float)
    lib.rtbh_scene_create.argtypes = [C.c_int, C.c_ulonglong, C.c_int, C.c_int, C.c_int, C.c_int]
    lib.rtbh_scene_create.restype = C.c_void_p
    lib.rtbh_scene_counts.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    lib.rtbh_scene_export.argtypes = [C.c_void_p, fp, fp, fp, fp, fp]
    lib.rtbh_scene_export_recipes.argtypes = [C.c_void_p, fp]
    lib.rtbh_scene_free.argtypes = [C.c_void_p]
    lib.rtbh_cube_build.argtypes = [fp, C.c_int, C.POINTER(C.c_int), fp, fp]
    lib.rtbh_convert_angle_to_radian.argtypes = [C.c_float]
    lib.rtbh_convert_angle_to_radian.restype = C.c_float
    lib.rtbh_orbit_camera.argtypes = [C.c_int, C.c_int, C.c_int, fp]
    lib.rtbh_png_write.argtypes = [C.c_char_p, C.c_void_p, C.c_uint, C.c_uint]
    lib.rtbh_png_write.restype = C.c_uint
    _lib = lib
    return lib


def _fp(a):
    return a.ctypes.data_as(C.POINTER(C.c_float))


class SceneArrays:
    """Flattened scene in the OpenCL kernel's argument layout (rayTracer.cl:111-114)."""

    def __init__(self, sph_o, sph_r, sph_c, cube_v, cube_c, label=""):
        self.sph_o, self.sph_r, self.sph_c = sph_o, sph_r, sph_c
        self.cube_v, self.cube_c = cube_v, cube_c
        self.label = label
        self.cube_recipes = None  # [nc, 7] (scale, rz, ry, rx, tx, ty, tz) for generated cubes

    @property
    def ns(self):
        return len(self.sph_r)

    @property
    def nc(self):
        return len(self.cube_c)

    def as_tuple(self):
        return self.sph_o, self.sph_r, self.sph_c, self.cube_v, self.cube_c


def _build(kind, seed, width, height, ns, nc, label):
    lib = load_host_library()
    h = lib.rtbh_scene_create(kind, seed, width, height, ns, nc)
    if not h:
        raise ValueError("bad scene request %r" % ((kind, seed, width, height, ns, nc),))
    try:
        a, b = C.c_int(), C.c_int()
        lib.rtbh_scene_counts(h, C.byref(a), C.byref(b))
        ns_, nc_ = a.value, b.value
        so, sr, sc = np.zeros((ns_, 4), np.float32), np.zeros(ns_, np.float32), np.zeros((ns_, 4), np.float32)
        cv, cc = np.zeros((nc_, 36, 4), np.float32), np.zeros((nc_, 4), np.float32)
        lib.rtbh_scene_export(h, _fp(so), _fp(sr), _fp(sc), _fp(cv), _fp(cc))
        nrec = lib.rtbh_scene_export_recipes(h, None)
        rec = np.zeros((nrec, 7), np.float32)
        if nrec:
            lib.rtbh_scene_export_recipes(h, _fp(rec))
    finally:
        lib.rtbh_scene_free(h)
    out = SceneArrays(so, sr, sc, cv, cc, label)
    out.cube_recipes = rec if nrec == nc_ else None
    return out


def reference_scene(which, seed=1):
    """createScene1 / createScene2 / createScene3 (RayTrace/states/MainState.cpp:419-639)."""
    return _build(int(which), seed, 0, 0, 0, 0, "scene%d(seed=%d)" % (which, seed))


def random_scene(width, height, ns, nc, seed):
    """createScene3's recipe spread over a width x height frame."""
    return _build(0, seed, width, height, ns, nc, "random(%dx%d,%d+%d,seed=%d)" % (width, height, ns, nc, seed))


def cube_from_ops(colour4, ops):
    lib = load_host_library()
    kinds = np.array([{"scale": 0, "rotate": 1, "translate": 2}[k] for k, _ in ops], np.int32)
    vecs = np.ascontiguousarray([v for _, v in ops], np.float32).reshape(-1, 3)
    col = np.ascontiguousarray(colour4, np.float32)
    out = np.zeros((36, 4), np.float32)
    lib.rtbh_cube_build(_fp(col), len(ops), kinds.ctypes.data_as(C.POINTER(C.c_int)), _fp(vecs), _fp(out))
    return out


def convert_angle_to_radian(a):
    return float(load_host_library().rtbh_convert_angle_to_radian(a))


def orbit_camera(width, height, k):
    """Camera k of the orbit sequence: (origin0, dU, dV, rayDir) rotated 3k degrees about y."""
    out = np.zeros(16, np.float32)
    load_host_library().rtbh_orbit_camera(width, height, k, _fp(out))
    return out[0:4].copy(), out[4:8].copy(), out[8:12].copy(), out[12:16].copy()


def png_write(path, rgba8):
    a = np.ascontiguousarray(rgba8, np.uint8)
    h, w = a.shape[0], a.shape[1]
    err = load_host_library().rtbh_png_write(path.encode(), a.ctypes.data, w, h)
    if err:
        raise IOError("encoder error %d" % err)
