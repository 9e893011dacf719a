"""GPU parity tests proper: the sm_100a kernels, called through the C ABI (include/rtb.h via
ctypes), against the oracle on the same inputs.

Bars (BASELINE.json north_star):
  * primitive-ID image identical on every non-silhouette pixel;
  * RGB within 1/255 per channel on >= 99.9 % of pixels, the rest only on silhouettes;
  * kernel == scalar emulator of the same header BIT FOR BIT (integer outputs), so every CPU
    emulator-vs-oracle result in test_kernel_math_emulated.py transfers to the device;
  * known-answer cases reproduce the reference's own ints exactly.
"""
import os

import numpy as np
import pytest

from parity import GOLDEN, assert_parity, kat_names, load_npz_scene, parity_report, to_oracle_scene

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True, params=["dense", "auto"])
def render_mode(request, gpu_ctx):
    """Every test of this file runs on the dense O(rays x primitives) kernel (the roofline kernel) and in the library's
    default mode (RTB_OPT_TILE_CULL = auto: binned from 64 primitives on): both must meet the same bars."""
    from opencl_ray_tracer_b200 import capi
    gpu_ctx.set_option(capi.RTB_OPT_TILE_CULL, capi.TILE_CULL_DENSE if request.param == "dense" else capi.TILE_CULL_AUTO)
    yield request.param
    gpu_ctx.set_option(capi.RTB_OPT_TILE_CULL, capi.TILE_CULL_AUTO)


def _set(ctx, sc):
    ctx.set_scene(sc.sph_o, sc.sph_r, sc.sph_c, sc.cube_v, sc.cube_c)


def test_scene1_default_frame(gpu_ctx, port, emu):
    sc = load_npz_scene(np.load(os.path.join(GOLDEN, "scene1.npz")))
    _set(gpu_ctx, sc)
    g = gpu_ctx.render(640, 480)
    o = port.render(sc, 640, 480)
    rep = parity_report(o[0], o[1], g["rgba_i32"], g["prim_id"])
    assert_parity(rep)
    e = emu.render(sc, 640, 480)
    assert np.array_equal(g["rgba_i32"], e[0]) and np.array_equal(g["prim_id"], e[1])
    # the three output forms agree: rgba8 is the (uint8_t) narrowing of the int stream
    assert np.array_equal(g["rgba8"], (g["rgba_i32"] & 0xFF).astype(np.uint8))
    assert g["kernel_ms"] > 0 and g["total_ms"] >= g["kernel_ms"]


@pytest.mark.parametrize("name", kat_names())
def test_known_answer_cases_exact(gpu_ctx, name):
    z = np.load(os.path.join(GOLDEN, "kats.npz"))
    sc = load_npz_scene(z, name + "/")
    _set(gpu_ctx, sc)
    g = gpu_ctx.render(64, 48)
    assert np.array_equal(g["rgba_i32"], z[name + "/rgba"])


@pytest.mark.parametrize("w,h,ns,nc,seed", [
    (640, 480, 100, 100, 7),      # reference scene-3 size
    (1920, 1080, 64, 64, 2),      # BASELINE config 2
    (333, 201, 40, 3, 11),        # ragged: width % 4 != 0, height % 4 != 0 -> scalar store path
    (1, 1, 5, 5, 1), (7, 3, 5, 5, 1),
    (512, 256, 0, 64, 5), (512, 256, 64, 0, 6),
    (640, 360, 769, 257, 13),     # crosses the staged-chunk boundaries (768 spheres / 256 cubes)
    (400, 300, 1537, 513, 14),    # three chunks of each kind
])
def test_random_scenes_vs_oracle(gpu_ctx, port, emu, w, h, ns, nc, seed):
    from opencl_ray_tracer_b200 import host_scene as hs
    s = hs.random_scene(max(w, 64), max(h, 64), ns, nc, seed)
    sc = to_oracle_scene(s)
    _set(gpu_ctx, sc)
    g = gpu_ctx.render(w, h)
    o = port.render(sc, w, h)
    if w * h >= 4096:
        assert_parity(parity_report(o[0], o[1], g["rgba_i32"], g["prim_id"]))
    e = emu.render(sc, w, h)
    assert np.array_equal(g["rgba_i32"], e[0]) and np.array_equal(g["prim_id"], e[1])


def test_empty_scene_and_errors(gpu_ctx):
    from opencl_ray_tracer_b200 import capi
    z = np.zeros((0, 4), np.float32)
    gpu_ctx.set_scene(z, np.zeros(0, np.float32), z, np.zeros((0, 36, 4), np.float32), z)
    g = gpu_ctx.render(65, 33)
    assert (g["rgba_i32"] == [0, 0, 0, 255]).all() and (g["prim_id"] == -1).all()
    with pytest.raises(capi.RtbError) as ei:
        gpu_ctx.render(0, 10)
    assert ei.value.status == 1
    o0, du, dv, d = capi.REFERENCE_CAMERA
    with pytest.raises(capi.RtbError) as ei:
        gpu_ctx.render(16, 16, camera=(o0, du, du, d))
    assert ei.value.status == 6
    fresh = capi.Context(0)
    with pytest.raises(capi.RtbError) as ei:
        fresh.render(16, 16)
    assert ei.value.status == 4  # RTB_ERR_NO_SCENE
    fresh.close()


@pytest.mark.parametrize("k", [1, 45, 119])
def test_orbit_cameras(gpu_ctx, port, k):
    from opencl_ray_tracer_b200 import host_scene as hs
    s = hs.random_scene(480, 360, 48, 48, 9)
    sc = to_oracle_scene(s)
    _set(gpu_ctx, sc)
    cam = hs.orbit_camera(480, 360, k)
    g = gpu_ctx.render(480, 360, camera=cam)
    o = port.render(sc, 480, 360, *cam)
    assert_parity(parity_report(o[0], o[1], g["rgba_i32"], g["prim_id"]))


def test_explicit_ray_origins_entry_point(gpu_ctx, port):
    """rtb_render_rays: the OpenCL kernel's rayOrigins[] + rayDir form (rayTracer.cl:111-114)."""
    from opencl_ray_tracer_b200 import host_scene as hs
    w, h = 320, 200
    s = hs.random_scene(w, h, 30, 30, 21)
    sc = to_oracle_scene(s)
    _set(gpu_ctx, sc)
    origins = port.pixel_origins(w, h)
    d = np.array([0, 0, -1, -1], np.float32)
    g = gpu_ctx.render_rays(origins, d)
    o = port.render(sc, w, h, ray_origins=origins, ray_dir=d)
    assert_parity(parity_report(o[0], o[1], g["rgba_i32"].reshape(h, w, 4), g["prim_id"].reshape(h, w)))
    # a shuffled, ragged subset of rays gives the same per-ray answers (no grid assumption)
    idx = np.random.default_rng(0).permutation(w * h)[:12345]
    g2 = gpu_ctx.render_rays(origins[idx], d)
    assert np.array_equal(g2["rgba_i32"], g["rgba_i32"][idx]) and np.array_equal(g2["prim_id"], g["prim_id"][idx])
    g0 = gpu_ctx.render_rays(np.zeros((0, 4), np.float32), d)
    assert g0["prim_id"].size == 0


@pytest.mark.parametrize("w,h,ns,nc,seed", [(640, 480, 0, 0, 0), (640, 480, 100, 100, 7), (1920, 1080, 64, 64, 2)])
def test_parity_with_the_opencl_path(gpu_ctx, port, w, h, ns, nc, seed):
    """Second parity target: the reference's OpenCL kernel (float triangle test),
    through the port's path="opencl" restatement, which tests/test_oracle_cl_variant.py pins to the
    kernel's own compiled source.  Same bar as against the CPU path."""
    from opencl_ray_tracer_b200 import host_scene as hs
    if ns + nc == 0:
        sc = load_npz_scene(np.load(os.path.join(GOLDEN, "scene1.npz")))
    else:
        sc = to_oracle_scene(hs.random_scene(w, h, ns, nc, seed))
    _set(gpu_ctx, sc)
    g = gpu_ctx.render(w, h)
    o = port.render(sc, w, h, path="opencl")
    rep = parity_report(o[0], o[1], g["rgba_i32"], g["prim_id"])
    assert_parity(rep)
