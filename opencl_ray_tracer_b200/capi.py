"""ctypes binding of the C ABI in include/rtb.h (librtb.so, built in-tree by build.sh).

This is plumbing only: every render goes to the hand-written sm_100a kernels inside librtb.so.
There is no fallback -- if the library is missing or no B200-class device is present the calls
raise.  Nothing in this package imports oracle/.
"""
import ctypes as C
import os

import numpy as np

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(PKG_DIR, "librtb.so")

RTB_OK = 0
STATUS_NAMES = {
    0: "RTB_OK", 1: "RTB_ERR_INVALID_ARG", 2: "RTB_ERR_NO_DEVICE", 3: "RTB_ERR_CUDA",
    4: "RTB_ERR_NO_SCENE", 5: "RTB_ERR_NOT_A_BOX", 6: "RTB_ERR_CAMERA",
    7: "RTB_ERR_OUT_OF_MEMORY", 8: "RTB_ERR_UNSUPPORTED",
}

# every symbol include/rtb.h declares (tests check that the library exports all of them)
EXPORTED_SYMBOLS = [
    "rtb_init", "rtb_shutdown", "rtb_error_string", "rtb_last_error_message", "rtb_set_scene",
    "rtb_render", "rtb_render_rays", "rtb_render_bands_device", "rtb_band_local_rows",
    "rtb_render_sequence_device", "rtb_host_alloc", "rtb_host_free", "rtb_device_info",
    "rtb_launch_count", "rtb_kernel_ms_avg", "rtb_debug_pack", "rtb_debug_pack_bounds", "rtb_debug_grid_shape", "rtb_version",
    "rtb_assemble_bands_device", "rtb_transfer_bytes", "rtb_render_bands_frame", "rtb_peer_frame_create",
    "rtb_peer_frame_open", "rtb_peer_frame_close", "rtb_debug_device_records", "rtb_render_bands_host",
    "rtb_host_register", "rtb_host_unregister", "rtb_render_bands_frame_signal", "rtb_stream_wait_signal",
    "rtb_peer_frame_signal_offset", "rtb_set_option", "rtb_get_option", "rtb_prepack_cameras",
    "rtb_prepack_cameras_bands", "rtb_debug_pack_row_table", "rtb_debug_device_row_table",
    "rtb_group_join", "rtb_group_leave", "rtb_group_info", "rtb_group_last_error", "rtb_group_render",
    "rtb_group_render_host", "rtb_group_prepack", "rtb_group_barrier", "rtb_render_device_noalloc",
]

RTB_OPT_TILE_CULL = 1
RTB_OPT_DEFER_SCENE_CHECK = 2
TILE_CULL_DENSE, TILE_CULL_PER_CTA, TILE_CULL_BINNED, TILE_CULL_AUTO = 0, 1, 2, 3
RTB_MAX_GROUP = 16


class RtbError(RuntimeError):
    def __init__(self, status, detail=""):
        self.status = status
        name = STATUS_NAMES.get(status, "RTB_ERR_%d" % status)
        super().__init__("%s%s" % (name, (": " + detail) if detail else ""))


_lib = None


def load_library(path=None):
    """Loads librtb.so; raises (loudly) when the CUDA extension has not been built."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = path or os.environ.get("RTB_LIB") or LIB_PATH   # RTB_LIB: tools/ A/B runs against an experiments build
    if not os.path.isfile(p):
        raise RuntimeError("librtb.so not found at %s -- run opencl_ray_tracer_b200/build.sh "
                           "(or __graft_entry__.build()); there is no CPU fallback" % p)
    lib = C.CDLL(p)
    fp, ip, vp = C.POINTER(C.c_float), C.POINTER(C.c_int32), C.c_void_p
    lib.rtb_init.argtypes = [C.POINTER(vp), C.c_int, C.POINTER(C.c_int)]
    lib.rtb_shutdown.argtypes = [vp]
    lib.rtb_shutdown.restype = None
    lib.rtb_error_string.argtypes = [C.c_int]
    lib.rtb_error_string.restype = C.c_char_p
    lib.rtb_last_error_message.argtypes = [vp]
    lib.rtb_last_error_message.restype = C.c_char_p
    lib.rtb_set_scene.argtypes = [vp, C.c_int, fp, fp, fp, C.c_int, fp, fp]
    lib.rtb_render.argtypes = [vp, C.c_int, C.c_int, fp, fp, fp, fp, vp, vp, vp, fp, fp]
    lib.rtb_render_rays.argtypes = [vp, C.c_size_t, fp, fp, vp, vp, vp, fp, fp]
    lib.rtb_render_bands_device.argtypes = [vp, C.c_int, C.c_int, fp, fp, fp, fp, C.c_int, C.c_int,
                                            C.c_int, vp, vp, vp, vp]
    lib.rtb_band_local_rows.argtypes = [C.c_int] * 4
    lib.rtb_render_bands_frame.argtypes = [vp, C.c_int, C.c_int, fp, fp, fp, fp, C.c_int, C.c_int, C.c_int, vp, vp]
    lib.rtb_render_bands_host.argtypes = [vp, C.c_int, C.c_int, fp, fp, fp, fp, C.c_int, C.c_int, C.c_int, vp, fp]
    lib.rtb_host_register.argtypes = [vp, C.c_size_t]
    lib.rtb_host_unregister.argtypes = [vp]
    lib.rtb_render_bands_frame_signal.argtypes = [vp, C.c_int, C.c_int, fp, fp, fp, fp, C.c_int, C.c_int, C.c_int, vp, vp, vp]
    lib.rtb_stream_wait_signal.argtypes = [vp, vp, C.c_uint, vp]
    lib.rtb_prepack_cameras.argtypes = [vp, C.c_int, C.c_int, C.c_int, fp, vp]
    lib.rtb_set_option.argtypes = [vp, C.c_int, C.c_int]
    lib.rtb_get_option.argtypes = [vp, C.c_int, C.POINTER(C.c_int)]
    lib.rtb_peer_frame_signal_offset.argtypes = [C.c_size_t]
    lib.rtb_peer_frame_signal_offset.restype = C.c_size_t
    lib.rtb_peer_frame_create.argtypes = [vp, C.c_size_t, C.POINTER(vp), C.POINTER(C.c_ubyte)]
    lib.rtb_peer_frame_open.argtypes = [vp, C.POINTER(C.c_ubyte), C.POINTER(vp)]
    lib.rtb_peer_frame_close.argtypes = [vp, vp, C.c_int]
    lib.rtb_assemble_bands_device.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, vp, vp, vp]
    lib.rtb_transfer_bytes.argtypes = [vp, C.POINTER(C.c_ulonglong), C.POINTER(C.c_ulonglong)]
    lib.rtb_transfer_bytes.restype = None
    lib.rtb_render_sequence_device.argtypes = [vp, C.c_int, C.c_int, C.c_int, fp, vp, fp]
    lib.rtb_host_alloc.argtypes = [C.POINTER(vp), C.c_size_t]
    lib.rtb_host_free.argtypes = [vp]
    lib.rtb_host_free.restype = None
    lib.rtb_device_info.argtypes = [vp, C.POINTER(C.c_int), C.POINTER(C.c_int), C.c_char_p, C.c_size_t]
    lib.rtb_launch_count.argtypes = [vp]
    lib.rtb_launch_count.restype = C.c_ulonglong
    lib.rtb_kernel_ms_avg.argtypes = [vp, C.c_int, C.POINTER(C.c_int)]
    lib.rtb_kernel_ms_avg.restype = C.c_float
    lib.rtb_debug_pack.argtypes = [C.c_int, fp, fp, C.c_int, fp, fp, fp, fp, fp, fp, fp, fp,
                                   C.POINTER(C.c_int)]
    lib.rtb_debug_pack_bounds.argtypes = [C.c_int, fp, fp, C.c_int, fp, C.c_int, C.c_int, fp, fp, fp, fp, fp]
    lib.rtb_debug_device_records.argtypes = [vp, C.c_int, C.c_int, fp, fp, fp, fp, fp, C.c_size_t,
                                             C.POINTER(C.c_size_t), C.POINTER(C.c_size_t)]
    lib.rtb_version.restype = C.c_char_p
    lib.rtb_prepack_cameras_bands.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, fp, vp]
    lib.rtb_debug_pack_row_table.argtypes = [C.c_int, fp, C.c_int, C.c_int, fp, fp, fp, fp, fp]
    lib.rtb_debug_device_row_table.argtypes = [vp, C.c_int, C.c_int, fp, fp, fp, fp, C.c_int, C.c_int, C.c_int, fp,
                                               C.c_size_t, C.POINTER(C.c_int)]
    lib.rtb_group_join.argtypes = [vp, C.c_char_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(vp)]
    lib.rtb_group_leave.argtypes = [vp]
    lib.rtb_group_info.argtypes = [vp, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int),
                                   C.POINTER(C.c_size_t)]
    lib.rtb_group_last_error.argtypes = [vp]
    lib.rtb_group_last_error.restype = C.c_char_p
    lib.rtb_group_render.argtypes = [vp, fp, fp, fp, fp, vp, C.POINTER(vp)]
    lib.rtb_group_render_host.argtypes = [vp, fp, fp, fp, fp, C.POINTER(vp), fp]
    lib.rtb_group_prepack.argtypes = [vp, C.c_int, fp, vp]
    lib.rtb_group_barrier.argtypes = [vp]
    lib.rtb_render_device_noalloc.argtypes = [C.c_int, C.c_int, C.c_int, fp, fp, fp, fp, C.c_int, vp, vp, vp, C.c_int, vp, vp,
                                              vp, vp, vp, vp, vp, C.c_char_p, C.c_size_t]
    if path is None:
        _lib = lib
    return lib


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _fp(a):
    return a.ctypes.data_as(C.POINTER(C.c_float)) if a is not None else None


def _vp(a):
    return C.c_void_p(a.ctypes.data) if a is not None else None


RTB_NOALLOC_MAX_PRIMITIVES = 1024
RTB_NOALLOC_MAX_BIN_WORDS = 131072


def render_device_noalloc(width, height, camera, ns, d_sph_o, d_sph_r, d_sph_c, nc, d_cube_v, d_cube_c, d_rgba8=0,
                          d_prim_id=0, d_rgba_i32=0, d_bad_cube=0, stream=0, device=-1):
    """rtb_render_device_noalloc: every d_* is a device address as an int (torch .data_ptr()); nothing is allocated
    by the call.  stream = cudaStream_t as int (0: default stream, synchronous)."""
    lib = load_library()
    o0, du, dv, d = [_f32(v) for v in camera]
    err = C.create_string_buffer(512)
    vpn = lambda a: C.c_void_p(a or None)
    rc = lib.rtb_render_device_noalloc(device, width, height, _fp(o0), _fp(du), _fp(dv), _fp(d), ns, vpn(d_sph_o),
                                       vpn(d_sph_r), vpn(d_sph_c), nc, vpn(d_cube_v), vpn(d_cube_c), vpn(d_rgba8),
                                       vpn(d_prim_id), vpn(d_rgba_i32), vpn(d_bad_cube), vpn(stream), err, len(err))
    if rc != RTB_OK:
        raise RtbError(rc, err.value.decode(errors="replace"))


# the reference's ray set-up, RayTrace/states/MainState.cpp:37-50
REFERENCE_CAMERA = (
    np.array([0, 0, 0, 1], np.float32),   # origin0
    np.array([1, 0, 0, 0], np.float32),   # dU
    np.array([0, 1, 0, 0], np.float32),   # dV
    np.array([0, 0, -1, -1], np.float32),  # rayDir = perspective(45,4/3,0,100)*(0,0,1,1)
)


def debug_pack(sph_o, sph_r, cube_v, camera=REFERENCE_CAMERA):
    """Host-side packer only (no device): returns (cube_rec[nc,3,4], sphere_rec[ns,4], cam[8])."""
    lib = load_library()
    sph_o, sph_r, cube_v = _f32(sph_o).reshape(-1, 4), _f32(sph_r).reshape(-1), _f32(cube_v).reshape(-1, 36, 4)
    ns, nc = len(sph_r), len(cube_v)
    cube_rec = np.zeros((nc, 3, 4), np.float32)
    sph_rec = np.zeros((ns, 4), np.float32)
    cam = np.zeros(8, np.float32)
    bad = C.c_int(-1)
    o0, du, dv, d = [_f32(v) for v in camera]
    rc = lib.rtb_debug_pack(ns, _fp(sph_o), _fp(sph_r), nc, _fp(cube_v), _fp(o0), _fp(du), _fp(dv),
                            _fp(d), _fp(cube_rec), _fp(sph_rec), _fp(cam), C.byref(bad))
    if rc != RTB_OK:
        raise RtbError(rc, "cube %d" % bad.value if bad.value >= 0 else "")
    return cube_rec, sph_rec, cam


def debug_grid_shape(width, rows, sm_count):
    """-> (rays per thread, tile width, tile height, CTAs) of the grid kernel for this launch."""
    out = (C.c_int * 4)()
    rc = load_library().rtb_debug_grid_shape(int(width), int(rows), int(sm_count), out)
    if rc != RTB_OK:
        raise RtbError(rc)
    return tuple(out)


def debug_pack_bounds(sph_o, sph_r, cube_v, width, height, camera=REFERENCE_CAMERA):
    """Host-side packer only: the conic bound records [nc+ns, 8] of a width x height frame."""
    lib = load_library()
    sph_o, sph_r, cube_v = _f32(sph_o).reshape(-1, 4), _f32(sph_r).reshape(-1), _f32(cube_v).reshape(-1, 36, 4)
    ns, nc = len(sph_r), len(cube_v)
    rec = np.zeros((nc + ns, 8), np.float32)
    o0, du, dv, d = [_f32(v) for v in camera]
    rc = lib.rtb_debug_pack_bounds(ns, _fp(sph_o), _fp(sph_r), nc, _fp(cube_v), int(width), int(height),
                                   _fp(o0), _fp(du), _fp(dv), _fp(d), _fp(rec))
    if rc != RTB_OK:
        raise RtbError(rc)
    return rec


def debug_pack_row_table(bound_rec, width, height, camera=REFERENCE_CAMERA):
    """Host-side packer only: the dense kernel's row table [height, n_bound, 2] for bound records [n_bound, 8]."""
    lib = load_library()
    b = _f32(bound_rec).reshape(-1, 8)
    tab = np.zeros((height, len(b), 2), np.float32)
    o0, du, dv, d = [_f32(v) for v in camera]
    rc = lib.rtb_debug_pack_row_table(len(b), _fp(b), int(width), int(height), _fp(o0), _fp(du), _fp(dv), _fp(d), _fp(tab))
    if rc != RTB_OK:
        raise RtbError(rc)
    return tab


class Context:
    """One rtb_ctx: one device (the one-process-per-GPU form), or -- devices=[...] -- several GPUs of this box
    driven from this process (rtb_init with n_devices > 1)."""

    def __init__(self, device=0, devices=None):
        self.lib = load_library()
        self._h = C.c_void_p()
        ids = list(devices) if devices is not None else [device]
        dev = (C.c_int * len(ids))(*ids)
        rc = self.lib.rtb_init(C.byref(self._h), len(ids), dev)
        if rc != RTB_OK:
            self._h = C.c_void_p()
            raise RtbError(rc, "rtb_init(devices=%s)" % ids)
        self.device = ids[0]
        self.n_devices = len(ids)
        self.ns = self.nc = 0

    def _check(self, rc):
        if rc != RTB_OK:
            raise RtbError(rc, (self.lib.rtb_last_error_message(self._h) or b"").decode())

    def close(self):
        if self._h:
            self.lib.rtb_shutdown(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def prepack_cameras(self, width, height, cameras, stream=None):
        """rtb_prepack_cameras: `cameras` = array (n, 16) of origin0, dU, dV, rayDir per frame."""
        cams = np.ascontiguousarray(cameras, np.float32).reshape(-1, 16)
        self._check(self.lib.rtb_prepack_cameras(self._h, int(width), int(height), len(cams), _fp(cams), C.c_void_p(stream or 0)))

    def prepack_cameras_bands(self, width, height, band_rows, rank, world, cameras, stream=None):
        cams = np.ascontiguousarray(cameras, np.float32).reshape(-1, 16)
        self._check(self.lib.rtb_prepack_cameras_bands(self._h, int(width), int(height), int(band_rows), int(rank), int(world),
                                                       len(cams), _fp(cams), C.c_void_p(stream or 0)))

    def debug_device_row_table(self, width, height, camera=REFERENCE_CAMERA, band_rows=16, rank=0, world=1):
        """The dense kernel's row table as the DEVICE builds it: [height, n_bound, 2] (rows of other ranks' bands 0)."""
        o0, du, dv, d = [_f32(v) for v in camera]
        n = C.c_int()
        self._check(self.lib.rtb_debug_device_row_table(self._h, width, height, _fp(o0), _fp(du), _fp(dv), _fp(d), band_rows,
                                                        rank, world, None, 0, C.byref(n)))
        tab = np.zeros((height, n.value, 2), np.float32)
        self._check(self.lib.rtb_debug_device_row_table(self._h, width, height, _fp(o0), _fp(du), _fp(dv), _fp(d), band_rows,
                                                        rank, world, _fp(tab), tab.size, C.byref(n)))
        return tab

    def set_option(self, option, value):
        """rtb_set_option, e.g. set_option(RTB_OPT_TILE_CULL, 1)."""
        self._check(self.lib.rtb_set_option(self._h, int(option), int(value)))

    def get_option(self, option):
        v = C.c_int()
        self._check(self.lib.rtb_get_option(self._h, int(option), C.byref(v)))
        return v.value

    def device_info(self):
        sm, khz = C.c_int(), C.c_int()
        name = C.create_string_buffer(128)
        self._check(self.lib.rtb_device_info(self._h, C.byref(sm), C.byref(khz), name, 128))
        return {"sm_count": sm.value, "sm_clock_khz": khz.value, "name": name.value.decode()}

    def set_scene(self, sph_o, sph_r, sph_c, cube_v, cube_c):
        sph_o, sph_r, sph_c = _f32(sph_o).reshape(-1, 4), _f32(sph_r).reshape(-1), _f32(sph_c).reshape(-1, 4)
        cube_v, cube_c = _f32(cube_v).reshape(-1, 36, 4), _f32(cube_c).reshape(-1, 4)
        self._check(self.lib.rtb_set_scene(self._h, len(sph_r), _fp(sph_o), _fp(sph_r), _fp(sph_c),
                                           len(cube_c), _fp(cube_v), _fp(cube_c)))
        self.ns, self.nc = len(sph_r), len(cube_c)

    def render(self, width, height, camera=REFERENCE_CAMERA, want=("rgba8", "prim_id", "rgba_i32"),
               out=None):
        """Blocking host-buffer render (rtb_render).  Returns dict with the requested arrays plus
        kernel_ms / total_ms.  `out` may hold preallocated arrays (e.g. pinned) keyed like `want`."""
        o0, du, dv, d = [_f32(v) for v in camera]
        out = dict(out or {})
        if "rgba8" in want and "rgba8" not in out:
            out["rgba8"] = np.empty((height, width, 4), np.uint8)
        if "prim_id" in want and "prim_id" not in out:
            out["prim_id"] = np.empty((height, width), np.int32)
        if "rgba_i32" in want and "rgba_i32" not in out:
            out["rgba_i32"] = np.empty((height, width, 4), np.int32)
        kms, tms = C.c_float(), C.c_float()
        self._check(self.lib.rtb_render(
            self._h, width, height, _fp(o0), _fp(du), _fp(dv), _fp(d),
            _vp(out.get("rgba8") if "rgba8" in want else None),
            _vp(out.get("prim_id") if "prim_id" in want else None),
            _vp(out.get("rgba_i32") if "rgba_i32" in want else None), C.byref(kms), C.byref(tms)))
        out["kernel_ms"], out["total_ms"] = kms.value, tms.value
        return out

    def render_rays(self, origins, ray_dir, want=("rgba8", "prim_id", "rgba_i32")):
        o = _f32(origins).reshape(-1, 4)
        d = _f32(ray_dir)
        n = len(o)
        out = {}
        if "rgba8" in want:
            out["rgba8"] = np.empty((n, 4), np.uint8)
        if "prim_id" in want:
            out["prim_id"] = np.empty(n, np.int32)
        if "rgba_i32" in want:
            out["rgba_i32"] = np.empty((n, 4), np.int32)
        kms, tms = C.c_float(), C.c_float()
        self._check(self.lib.rtb_render_rays(self._h, n, _fp(o), _fp(d), _vp(out.get("rgba8")),
                                             _vp(out.get("prim_id")), _vp(out.get("rgba_i32")),
                                             C.byref(kms), C.byref(tms)))
        out["kernel_ms"], out["total_ms"] = kms.value, tms.value
        return out

    def band_local_rows(self, height, band_rows, rank, world):
        return self.lib.rtb_band_local_rows(height, band_rows, rank, world)

    def render_bands_device(self, width, height, camera, band_rows, rank, world, d_rgba8=0,
                            d_prim_id=0, d_rgba_i32=0, stream=0):
        """Device-pointer render (ints from torch .data_ptr()); stream = cudaStream_t as int
        (0 -> the context's own stream, synchronous)."""
        o0, du, dv, d = [_f32(v) for v in camera]
        self._check(self.lib.rtb_render_bands_device(
            self._h, width, height, _fp(o0), _fp(du), _fp(dv), _fp(d), band_rows, rank, world,
            C.c_void_p(d_rgba8 or None), C.c_void_p(d_prim_id or None), C.c_void_p(d_rgba_i32 or None),
            C.c_void_p(stream or None)))

    def render_bands_frame(self, width, height, camera, band_rows, rank, world, d_frame_rgba8, stream=0, d_signal=0):
        """This rank's bands written straight into a whole frame (own or peer-mapped memory); with
        d_signal the launch's last CTA adds 1 to that word when all of its stores are ordered."""
        o0, du, dv, d = [_f32(v) for v in camera]
        self._check(self.lib.rtb_render_bands_frame_signal(
            self._h, width, height, _fp(o0), _fp(du), _fp(dv), _fp(d), band_rows, rank, world,
            C.c_void_p(d_frame_rgba8), C.c_void_p(d_signal or None), C.c_void_p(stream or None)))

    def stream_wait_signal(self, d_signal, value, stream=0):
        """Makes `stream` wait until the 32-bit word at d_signal has reached `value` (wrap-safe >=)."""
        self._check(self.lib.rtb_stream_wait_signal(self._h, C.c_void_p(d_signal), C.c_uint(value & 0xFFFFFFFF),
                                                    C.c_void_p(stream or None)))

    def render_bands_host(self, width, height, camera, band_rows, rank, world, host_frame):
        """This rank's bands into a whole HOST frame (numpy uint8 [H, W, 4], e.g. over shared memory);
        blocking.  Returns the grid kernel's device time in ms."""
        o0, du, dv, d = [_f32(v) for v in camera]
        kms = C.c_float()
        self._check(self.lib.rtb_render_bands_host(
            self._h, width, height, _fp(o0), _fp(du), _fp(dv), _fp(d), band_rows, rank, world,
            C.c_void_p(host_frame.ctypes.data), C.byref(kms)))
        return kms.value

    def peer_frame_create(self, nbytes):
        """-> (device pointer, 64-byte IPC handle as bytes)."""
        p = C.c_void_p()
        h = (C.c_ubyte * 64)()
        self._check(self.lib.rtb_peer_frame_create(self._h, C.c_size_t(nbytes), C.byref(p), h))
        return p.value, bytes(h)

    def peer_frame_open(self, handle):
        p = C.c_void_p()
        h = (C.c_ubyte * 64).from_buffer_copy(handle)
        self._check(self.lib.rtb_peer_frame_open(self._h, h, C.byref(p)))
        return p.value

    def peer_frame_close(self, ptr, owner):
        self._check(self.lib.rtb_peer_frame_close(self._h, C.c_void_p(ptr), int(bool(owner))))

    def assemble_bands_device(self, width, height, band_rows, world, elem_bytes, d_packed, d_frame, stream=0):
        self._check(self.lib.rtb_assemble_bands_device(self._h, width, height, band_rows, world, elem_bytes,
                                                       C.c_void_p(d_packed), C.c_void_p(d_frame),
                                                       C.c_void_p(stream or None)))

    def transfer_bytes(self):
        a, b = C.c_ulonglong(), C.c_ulonglong()
        self.lib.rtb_transfer_bytes(self._h, C.byref(a), C.byref(b))
        return a.value, b.value

    def render_sequence_device(self, width, height, cameras, d_rgba8):
        cams = _f32(cameras).reshape(-1, 16)
        kms = C.c_float()
        self._check(self.lib.rtb_render_sequence_device(self._h, width, height, len(cams), _fp(cams),
                                                        C.c_void_p(d_rgba8), C.byref(kms)))
        return kms.value

    def debug_device_records(self, width, height, camera=REFERENCE_CAMERA):
        """The frame records as the DEVICE pack kernel writes them:
        (cube_rec[nc,3,4], sphere_rec[ns,4], bound_rec[n_bound,8])."""
        o0, du, dv, d = [_f32(v) for v in camera]
        n, off = C.c_size_t(), C.c_size_t()
        self._check(self.lib.rtb_debug_device_records(self._h, width, height, _fp(o0), _fp(du), _fp(dv), _fp(d),
                                                      None, 0, C.byref(n), C.byref(off)))
        buf = np.zeros(n.value, np.float32)
        self._check(self.lib.rtb_debug_device_records(self._h, width, height, _fp(o0), _fp(du), _fp(dv), _fp(d),
                                                      _fp(buf), buf.size, C.byref(n), C.byref(off)))
        nc, ns = self.nc, self.ns
        return (buf[:12 * nc].reshape(nc, 3, 4).copy(), buf[12 * nc:12 * nc + 4 * ns].reshape(ns, 4).copy(),
                buf[off.value:].reshape(-1, 8).copy())

    def launch_count(self):
        return int(self.lib.rtb_launch_count(self._h))

    def kernel_ms_avg(self, reset=True):
        n = C.c_int()
        avg = self.lib.rtb_kernel_ms_avg(self._h, 1 if reset else 0, C.byref(n))
        return float(avg), n.value


class Group:
    """rtb_group_*: this process's membership in a multi-GPU job (one process per GPU).  Every rank constructs it
    with the same name / sizes (collective) and calls render() / render_host() once per frame."""

    def __init__(self, ctx, name, rank, world, width, height, band_rows=16, slots=3):
        self.ctx, self.rank, self.world, self.width, self.height = ctx, rank, world, width, height
        self.lib = ctx.lib
        self._g = C.c_void_p()
        rc = self.lib.rtb_group_join(ctx._h, name.encode(), rank, world, width, height, band_rows, slots, C.byref(self._g))
        if rc != RTB_OK:
            self._g = C.c_void_p()
            raise RtbError(rc, (self.lib.rtb_last_error_message(ctx._h) or b"").decode())
        sb = C.c_size_t()
        self.lib.rtb_group_info(self._g, None, None, None, None, C.byref(sb))
        self.slot_bytes = sb.value

    def _check(self, rc):
        if rc != RTB_OK:
            raise RtbError(rc, (self.lib.rtb_group_last_error(self._g) or b"").decode() or
                           (self.lib.rtb_last_error_message(self.ctx._h) or b"").decode())

    def render(self, camera, stream=0):
        """Enqueues this rank's share of the next frame; rank 0 gets the device pointer of the frame's slot."""
        o0, du, dv, d = [_f32(v) for v in camera]
        p = C.c_void_p()
        self._check(self.lib.rtb_group_render(self._g, _fp(o0), _fp(du), _fp(dv), _fp(d), C.c_void_p(stream or None), C.byref(p)))
        return p.value

    def render_host(self, camera):
        """Blocking; rank 0: (numpy uint8 [H, W, 4] view of the shared host frame, kernel ms); others: (None, kernel ms)."""
        o0, du, dv, d = [_f32(v) for v in camera]
        p, kms = C.c_void_p(), C.c_float()
        self._check(self.lib.rtb_group_render_host(self._g, _fp(o0), _fp(du), _fp(dv), _fp(d), C.byref(p), C.byref(kms)))
        if not p.value:
            return None, kms.value
        n = 4 * self.width * self.height
        buf = (C.c_uint8 * n).from_address(p.value)
        return np.frombuffer(buf, dtype=np.uint8).reshape(self.height, self.width, 4), kms.value

    def prepack(self, cameras, stream=0):
        cams = np.ascontiguousarray(cameras, np.float32).reshape(-1, 16)
        self._check(self.lib.rtb_group_prepack(self._g, len(cams), _fp(cams), C.c_void_p(stream or None)))

    def barrier(self):
        self._check(self.lib.rtb_group_barrier(self._g))

    def frame_tensor(self, ptr):
        """torch uint8 view [H, W, 4] of a slot pointer returned by render() (rank 0)."""
        import torch

        class _Raw:
            pass
        raw = _Raw()
        raw.__cuda_array_interface__ = {"shape": (self.height, self.width, 4), "typestr": "|u1",
                                        "data": (int(ptr), False), "version": 3, "strides": None}
        return torch.as_tensor(raw, device="cuda")

    def close(self):
        if self._g:
            self.lib.rtb_group_leave(self._g)
            self._g = C.c_void_p()


def host_alloc(nbytes):
    """Pinned host buffer as a uint8 numpy array (freed with host_free)."""
    lib = load_library()
    p = C.c_void_p()
    rc = lib.rtb_host_alloc(C.byref(p), nbytes)
    if rc != RTB_OK:
        raise RtbError(rc, "rtb_host_alloc")
    buf = (C.c_uint8 * nbytes).from_address(p.value)
    arr = np.frombuffer(buf, dtype=np.uint8)
    arr_ptr = p.value
    return arr, arr_ptr


def peer_frame_signal_offset(nbytes):
    return int(load_library().rtb_peer_frame_signal_offset(C.c_size_t(nbytes)))


def host_register(arr):
    """Page-locks a caller-owned numpy buffer (shared memory, ...) for direct DMA."""
    rc = load_library().rtb_host_register(C.c_void_p(arr.ctypes.data), C.c_size_t(arr.nbytes))
    if rc != RTB_OK:
        raise RtbError(rc, "rtb_host_register")


def host_unregister(arr):
    load_library().rtb_host_unregister(C.c_void_p(arr.ctypes.data))


def host_free(ptr):
    load_library().rtb_host_free(C.c_void_p(ptr))
