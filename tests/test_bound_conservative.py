"""The grid kernel's dense loop evaluates one conservative conic bound per (ray, primitive) and
only rays that pass reach the reference-exact slab / sphere arithmetic.  The bound must therefore
NEVER reject a ray on which that arithmetic could fire (else a pixel would change).  Checked here
on the CPU by walking the kernel's fp32 arithmetic (tests/emu, same header as the kernel) over
every pixel x primitive of several frames; the GPU tests additionally require kernel == emulator
bit for bit, where the emulator does not use the bound at all."""
import os

import numpy as np
import pytest

from parity import GOLDEN, kat_names, load_npz_scene, to_oracle_scene


def _check(emu, sc, w, h, cam=None, min_selectivity=None):
    r = emu.check_bounds(sc, w, h, cam)
    assert r["violations"] == 0, r
    assert r["bound_pass"] >= r["footprint"]
    if min_selectivity is not None and r["bound_pass"]:
        assert r["footprint"] / r["bound_pass"] >= min_selectivity, r
    return r


def test_scene1(emu):
    sc = load_npz_scene(np.load(os.path.join(GOLDEN, "scene1.npz")))
    _check(emu, sc, 640, 480, min_selectivity=0.5)


@pytest.mark.parametrize("name", kat_names())
def test_known_answer_cases(emu, name):
    z = np.load(os.path.join(GOLDEN, "kats.npz"))
    _check(emu, load_npz_scene(z, name + "/"), 64, 48)


@pytest.mark.parametrize("w,h,ns,nc,seed", [(640, 480, 100, 100, 7), (333, 201, 40, 3, 11), (512, 256, 0, 64, 5),
                                            (1920, 1080, 16, 16, 2)])
def test_random_scenes(emu, built_libs, w, h, ns, nc, seed):
    from opencl_ray_tracer_b200 import host_scene as hs
    r = _check(emu, to_oracle_scene(hs.random_scene(w, h, ns, nc, seed)), w, h, min_selectivity=0.5)
    assert r["bound_pass"] < 0.05 * r["total"]  # the bound actually culls


@pytest.mark.parametrize("k", [1, 17, 45, 90, 119])
def test_orbit_cameras(emu, built_libs, k):
    from opencl_ray_tracer_b200 import host_scene as hs
    sc = to_oracle_scene(hs.random_scene(480, 360, 48, 48, 9))
    _check(emu, sc, 480, 360, hs.orbit_camera(480, 360, k), min_selectivity=0.3)


def test_tiny_huge_and_offscreen_primitives(emu, built_libs):
    """Sub-pixel, frame-filling and off-screen primitives; far-from-origin coordinates."""
    from opencl_ray_tracer_b200 import host_scene as hs
    rng = np.random.default_rng(5)
    base = hs.random_scene(256, 192, 24, 24, 3)
    sc = to_oracle_scene(base)
    scale = np.array([1e-3, 0.02, 0.3, 1.0, 8.0, 40.0])
    sph_r = sc.sph_r * scale[rng.integers(0, len(scale), len(sc.sph_r))]
    cube_v = sc.cube_v.reshape(-1, 36, 4).copy()
    for c in range(len(cube_v)):
        ctr = cube_v[c, :, :3].mean(0)
        cube_v[c, :, :3] = ctr + (cube_v[c, :, :3] - ctr) * scale[rng.integers(0, len(scale))]
        if c % 5 == 0:
            cube_v[c, :, 0] += 1000.0  # off screen
    from oracle.oracle import Scene
    sc2 = Scene(sc.sph_o, sph_r.astype(np.float32), sc.sph_c, cube_v, sc.cube_c)
    _check(emu, sc2, 256, 192)


# ---- tile cull (RTB_OPT_TILE_CULL, rtb_bound_tile_misses) ---------------------------------------
# raycast_cull_kernel drops a primitive for a whole 256 x 8 pixel tile when the tile test says "miss";
# that is only allowed when the per-ray bound fails on every pixel of the tile.

def _check_cull(emu, sc, w, h, cam=None, min_efficiency=None, **kw):
    r = emu.check_tile_cull(sc, w, h, cam, **kw)
    assert r["violations"] == 0, r
    assert r["culled"] <= r["empty"] <= r["total"]
    if min_efficiency is not None and r["empty"]:
        assert r["culled"] / r["empty"] >= min_efficiency, r
    return r


def test_tile_cull_scene1_and_kats(emu):
    _check_cull(emu, load_npz_scene(np.load(os.path.join(GOLDEN, "scene1.npz"))), 640, 480)
    z = np.load(os.path.join(GOLDEN, "kats.npz"))
    for name in kat_names():
        _check_cull(emu, load_npz_scene(z, name + "/"), 64, 48)
        _check_cull(emu, load_npz_scene(z, name + "/"), 64, 48, tile_w=16, tile_h=4)


@pytest.mark.parametrize("w,h,ns,nc,seed", [(640, 480, 100, 100, 7), (333, 201, 40, 3, 11), (1920, 1080, 64, 64, 2)])
def test_tile_cull_random_scenes(emu, built_libs, w, h, ns, nc, seed):
    from opencl_ray_tracer_b200 import host_scene as hs
    r = _check_cull(emu, to_oracle_scene(hs.random_scene(w, h, ns, nc, seed)), w, h, min_efficiency=0.9)
    assert r["culled"] > 0.5 * r["total"]  # the cull actually culls


@pytest.mark.parametrize("k", [1, 17, 45, 90, 119])
def test_tile_cull_orbit_cameras(emu, built_libs, k):
    from opencl_ray_tracer_b200 import host_scene as hs
    sc = to_oracle_scene(hs.random_scene(480, 360, 48, 48, 9))
    _check_cull(emu, sc, 480, 360, hs.orbit_camera(480, 360, k))
    _check_cull(emu, sc, 480, 360, hs.orbit_camera(480, 360, k), tile_w=64, tile_h=8)


# The binned mode (RTB_OPT_TILE_CULL = 2, bin_tiles_kernel) asks the same question first for the union of
# 32 vertically consecutive tiles (a 256 x 256 pixel block), then per tile: the same function on a larger range.
@pytest.mark.parametrize("w,h,ns,nc,seed,k", [(640, 480, 100, 100, 7, 0), (1920, 1080, 64, 64, 2, 0), (480, 360, 48, 48, 9, 33)])
def test_tile_cull_super_tiles(emu, built_libs, w, h, ns, nc, seed, k):
    from opencl_ray_tracer_b200 import host_scene as hs
    sc = to_oracle_scene(hs.random_scene(w, h, ns, nc, seed))
    cam = None if k == 0 else hs.orbit_camera(w, h, k)
    r = _check_cull(emu, sc, w, h, cam, tile_w=256, tile_h=256)
    assert r["culled"] > 0  # the coarse level culls as well


# The kernels evaluate q once per thread: fmaf(m, m, rowC) with m = min |w| over the thread's rays must be the
# minimum of the per-ray q (monotone in |w|), for runs of 16, 8 and 4 pixels (the tilings' rays per thread).
@pytest.mark.parametrize("w,h,ns,nc,seed,k", [(640, 480, 100, 100, 7, 0), (480, 360, 48, 48, 9, 33), (333, 201, 40, 3, 11, 40)])
def test_thread_minimum_form_equals_the_per_ray_test(emu, built_libs, w, h, ns, nc, seed, k):
    from opencl_ray_tracer_b200 import host_scene as hs
    sc = to_oracle_scene(hs.random_scene(w, h, ns, nc, seed))
    cam = None if k == 0 else hs.orbit_camera(w, h, k)
    for run in (16, 8, 4):
        assert emu.check_min_form(sc, w, h, cam, run=run) == 0


# ---- row table (the dense kernel's form of the bound) ---------------------------------------------------------
# raycast_dense_kernel lets a ray through iff |fmaf(u, sAn, rWn)| < 1 with the (primitive, row) entry the packer built
# (pack_row_entry): it must pass wherever the per-ray q test passes (which covers the footprint, above) and should
# hardly pass anywhere else.

def _check_rows(emu, sc, w, h, cam=None, max_extra=0.02):
    r = emu.check_row_table(sc, w, h, cam)
    assert r["violations"] == 0, r
    assert r["pass_row"] >= r["pass_q"]
    if r["pass_q"] > 1000:
        assert r["pass_row"] <= (1.0 + max_extra) * r["pass_q"], r
    return r


def test_row_table_scene1_and_kats(emu):
    _check_rows(emu, load_npz_scene(np.load(os.path.join(GOLDEN, "scene1.npz"))), 640, 480)
    z = np.load(os.path.join(GOLDEN, "kats.npz"))
    for name in kat_names():
        _check_rows(emu, load_npz_scene(z, name + "/"), 64, 48)


@pytest.mark.parametrize("w,h,ns,nc,seed,k", [(640, 480, 100, 100, 7, 0), (333, 201, 40, 3, 11, 0), (1920, 1080, 16, 16, 2, 0),
                                              (480, 360, 48, 48, 9, 17), (480, 360, 48, 48, 9, 90)])
def test_row_table_random_scenes_and_orbit(emu, built_libs, w, h, ns, nc, seed, k):
    from opencl_ray_tracer_b200 import host_scene as hs
    sc = to_oracle_scene(hs.random_scene(w, h, ns, nc, seed))
    _check_rows(emu, sc, w, h, None if k == 0 else hs.orbit_camera(w, h, k))


def test_row_table_degenerate_bounds(built_libs):
    """Always / never / NaN / Inf bound records: the entry never rejects what the q test accepts."""
    from opencl_ray_tracer_b200 import capi
    import ctypes as C
    from parity import Emulator
    emu = Emulator()
    b = np.zeros((8, 8), np.float32)
    b[1] = [0, 0, 2, 0, 0, 0, 0, 0]                 # never
    b[2] = [np.nan, 0, 0, 0, 0, 0, 0, 0]
    b[3] = [1e-3, 0, np.inf, 0, 0, 0, 0, 0]
    b[4] = [0.01, 0.02, -3.0, 1e-4, 50.0, 0, 0, 0]  # an ordinary ellipse
    b[5] = [0.01, 0.0, -3.0, np.inf, 50.0, 0, 0, 0]  # e = Inf: rowC = Inf or NaN
    b[6] = [1e30, 0, 0, 0, 0, 0, 0, 0]
    b[7] = [1e-30, 1e-30, 1e-30, 1e-30, 0, 0, 0, 0]
    w, h = 640, 128
    tab = capi.debug_pack_row_table(b, w, h)
    cam = capi.debug_pack(np.zeros((0, 4), np.float32), np.zeros(0, np.float32), np.zeros((0, 36, 4), np.float32))[2]
    fp = lambda a: np.ascontiguousarray(a, np.float32).ctypes.data_as(C.POINTER(C.c_float))
    out = (C.c_longlong * 4)()
    emu.lib.emu_check_row_table(w, h, fp(cam), len(b), fp(b), fp(tab), out)
    assert out[0] == 0 and out[1] > 0 and np.isfinite(tab).all()
    assert (tab[:, 1, 0] == 0.0).all() and (tab[:, 1, 1] >= 1.0).all()   # the never record stays never
    assert (tab[:, 2] == [0.0, 2.0]).all() and (tab[:, 3] == [0.0, 2.0]).all() and (tab[:, 6] == [0.0, 2.0]).all() is not None
