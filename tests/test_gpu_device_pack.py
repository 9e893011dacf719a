"""The per-frame records are packed ON the device (scene_pack_device.cu, the shared code of
scene_pack_impl.h).  They must equal the host packer's records (rtb_debug_pack /
rtb_debug_pack_bounds) bit for bit, so that the CPU-side conservativeness and emulator tests
(test_bound_conservative.py, test_kernel_math_emulated.py) speak for what the kernels read."""
import math

import numpy as np
import pytest

from opencl_ray_tracer_b200 import capi, host_scene as hs

pytestmark = pytest.mark.gpu


def _orbit_camera(w, h, deg):
    a = math.radians(deg)
    c, s = math.cos(a), math.sin(a)
    ctr = np.array([w / 2, h / 2, -60.0])

    def rot(v):
        return np.array([c * v[0] + s * v[2], v[1], -s * v[0] + c * v[2]])
    o0 = ctr + rot(np.array([0, 0, 0.0]) - ctr)
    return (np.array([*o0, 1], np.float32), np.array([*rot([1, 0, 0]), 0], np.float32),
            np.array([*rot([0, 1, 0]), 0], np.float32), np.array([*rot([0, 0, -1]), -1], np.float32))


def _bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


@pytest.mark.parametrize("w,h,ns,nc,seed,deg", [
    (640, 480, 100, 100, 7, 0), (1920, 1080, 64, 64, 2, 0), (3840, 2160, 512, 512, 3, 0),
    (3840, 2160, 512, 512, 3, 33), (333, 201, 40, 3, 11, 120), (640, 360, 769, 257, 13, 0),
    (512, 256, 0, 64, 5, 3), (512, 256, 64, 0, 6, 357),
])
def test_device_records_equal_host_records(gpu_ctx, w, h, ns, nc, seed, deg):
    s = hs.random_scene(max(w, 64), max(h, 64), ns, nc, seed)
    cam = capi.REFERENCE_CAMERA if deg == 0 else _orbit_camera(w, h, deg)
    gpu_ctx.set_scene(*s.as_tuple())
    d_cube, d_sph, d_bound = gpu_ctx.debug_device_records(w, h, cam)
    h_cube, h_sph, _ = capi.debug_pack(s.sph_o, s.sph_r, s.cube_v, cam)
    h_bound = capi.debug_pack_bounds(s.sph_o, s.sph_r, s.cube_v, w, h, cam)
    assert np.array_equal(_bits(d_cube), _bits(h_cube))
    assert np.array_equal(_bits(d_sph), _bits(h_sph))
    n = nc + ns
    assert np.array_equal(_bits(d_bound[:n, :5]), _bits(h_bound[:, :5]))
    # padding up to the vote group: "never" records (w = 2, q = 4)
    assert len(d_bound) % 8 == 0 and len(d_bound) - n < 8
    assert np.all(d_bound[n:, 2] == 2.0) and np.all(d_bound[n:, [0, 1, 3, 4]] == 0.0)


def test_scene1_and_kat_records(gpu_ctx):
    s = hs.reference_scene(1)
    gpu_ctx.set_scene(*s.as_tuple())
    d_cube, d_sph, d_bound = gpu_ctx.debug_device_records(640, 480)
    h_cube, h_sph, _ = capi.debug_pack(s.sph_o, s.sph_r, s.cube_v)
    h_bound = capi.debug_pack_bounds(s.sph_o, s.sph_r, s.cube_v, 640, 480)
    assert np.array_equal(_bits(d_cube), _bits(h_cube)) and np.array_equal(_bits(d_sph), _bits(h_sph))
    assert np.array_equal(_bits(d_bound[:len(h_bound), :5]), _bits(h_bound[:, :5]))


def test_not_a_box_is_reported_from_the_device(gpu_ctx):
    s = hs.random_scene(640, 480, 4, 9, 21)
    cube_v = np.array(s.cube_v, np.float32).reshape(-1, 36, 4).copy()
    cube_v[5, 17, 1] += 3.0      # one vertex off its corner
    cube_v[7, 2, 0] -= 1.0
    with pytest.raises(capi.RtbError) as e:
        gpu_ctx.set_scene(s.sph_o, s.sph_r, s.sph_c, cube_v, s.cube_c)
    assert e.value.status == 5 and "cube 5" in str(e.value)
    with pytest.raises(capi.RtbError) as e2:   # the failed call left no scene behind
        gpu_ctx.render(64, 48)
    assert e2.value.status == 4
    gpu_ctx.set_scene(*s.as_tuple())           # and the context is still usable
    gpu_ctx.render(64, 48, want=("rgba8",))


@pytest.mark.parametrize("w,h,ns,nc,seed,deg,bands", [(640, 480, 100, 100, 7, 0, (16, 0, 1)), (333, 201, 40, 3, 11, 120, (16, 1, 3)),
                                                      (1920, 1080, 64, 64, 2, 45, (16, 5, 8)), (250, 40, 90, 13, 19, 21, (4, 0, 2))])
def test_device_row_table_equals_host_row_table(gpu_ctx, w, h, ns, nc, seed, deg, bands):
    """row_table_kernel (the dense kernel's per-(row, primitive) entries) against the host packer's, bit for bit, on the
    rows of the given bands; the CPU conservativeness walk (test_bound_conservative.py) speaks for the host's."""
    s = hs.random_scene(max(w, 64), max(h, 64), ns, nc, seed)
    cam = capi.REFERENCE_CAMERA if deg == 0 else _orbit_camera(w, h, deg)
    gpu_ctx.set_scene(*s.as_tuple())
    band_rows, rank, world = bands
    d_tab = gpu_ctx.debug_device_row_table(w, h, cam, band_rows, rank, world)
    _, _, d_bound = gpu_ctx.debug_device_records(w, h, cam)
    h_tab = capi.debug_pack_row_table(d_bound, w, h, cam)
    assert d_tab.shape == h_tab.shape and d_tab.shape[1] == len(d_bound)
    mine = np.array([(y // band_rows) % world == rank for y in range(h)])
    assert np.array_equal(_bits(d_tab[mine]), _bits(h_tab[mine]))
    assert not d_tab[~mine].any()                       # other ranks' rows are not built
    n = nc + ns
    assert np.all(d_tab[mine][:, n:, 0] == 0.0) and np.all(d_tab[mine][:, n:, 1] >= 1.0)   # padding never passes


def test_deferred_scene_check(built_libs):
    """RTB_OPT_DEFER_SCENE_CHECK: rtb_set_scene returns without waiting for the device; the NOT_A_BOX verdict comes
    from the next render call."""
    s = hs.random_scene(640, 480, 4, 9, 21)
    bad = np.array(s.cube_v, np.float32).reshape(-1, 36, 4).copy()
    bad[5, 17, 1] += 3.0
    with capi.Context(0) as c:
        c.set_option(capi.RTB_OPT_DEFER_SCENE_CHECK, 1)
        c.set_scene(s.sph_o, s.sph_r, s.sph_c, bad, s.cube_c)          # no error yet
        with pytest.raises(capi.RtbError) as e:
            c.render(64, 48, want=("rgba8",))
        assert e.value.status == 5 and "cube 5" in str(e.value)
        with pytest.raises(capi.RtbError) as e2:
            c.render(64, 48, want=("rgba8",))
        assert e2.value.status == 4                                     # no scene left behind
        c.set_scene(*s.as_tuple())
        a = c.render(64, 48, want=("rgba8",))["rgba8"]
        c.set_option(capi.RTB_OPT_DEFER_SCENE_CHECK, 0)
        c.set_scene(*s.as_tuple())
        assert np.array_equal(c.render(64, 48, want=("rgba8",))["rgba8"], a)
