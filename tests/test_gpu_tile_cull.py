"""RTB_OPT_TILE_CULL: raycast_cull_kernel (1) drops, per CTA tile, the primitives whose conic bound no
ray of the tile can pass (rtb_bound_tile_misses, a statement about the values the dense per-ray test
would compute); the binned mode (2) makes the same decision once per frame in bin_tiles_kernel and
raycast_binned_kernel renders from the per-tile bit words.  Both must therefore write exactly the bytes
of the dense kernel -- all three outputs, every entry point, any camera -- while being much faster on
SURVEY.md s8d's scenes."""
import numpy as np
import pytest

from opencl_ray_tracer_b200 import capi, host_scene as hs

pytestmark = pytest.mark.gpu


@pytest.fixture()
def ctx(gpu_ctx):
    yield gpu_ctx
    gpu_ctx.set_option(capi.RTB_OPT_TILE_CULL, capi.TILE_CULL_AUTO)


MODES = [1, 2, 3]   # per-CTA cull, binned, auto (the library's default)


def _both(ctx, w, h, cam, mode, want=("rgba8", "prim_id", "rgba_i32")):
    ctx.set_option(capi.RTB_OPT_TILE_CULL, 0)
    dense = ctx.render(w, h, cam, want=want)
    ctx.set_option(capi.RTB_OPT_TILE_CULL, mode)
    assert ctx.get_option(capi.RTB_OPT_TILE_CULL) == mode
    culled = ctx.render(w, h, cam, want=want)
    return dense, culled


@pytest.mark.parametrize("w,h,ns,nc,seed,k", [
    (640, 480, 100, 100, 7, 0), (1920, 1080, 64, 64, 2, 0), (3840, 2160, 512, 512, 3, 0),
    (3840, 2160, 512, 512, 3, 11), (3840, 2160, 512, 512, 3, 47), (333, 201, 40, 3, 11, 40),
    (640, 360, 769, 257, 13, 0), (512, 256, 0, 64, 5, 1), (512, 256, 64, 0, 6, 119),
    (128, 64, 400, 300, 17, 0),    # > 256 survivors per tile: the list is flushed mid-way
    (250, 40, 900, 131, 19, 7),
    (96, 40, 3000, 2000, 23, 0),   # 5000 primitives: more words than a CTA has threads, many passes
    (1000, 700, 1, 0, 29, 0), (64, 48, 0, 1, 31, 0),
    # dense launches with a short tail (more than 1.25 waves of CTAs: the last tile rows run in 8-ray CTAs on a second
    # stream) whose body and tail both end in partial tiles: width not a multiple of 4 / of 128, rows not a multiple of 4
    (2001, 1203, 90, 60, 37, 0), (2500, 1102, 60, 70, 41, 9), (1924, 1440, 100, 50, 43, 0),
])
@pytest.mark.parametrize("mode", MODES)
def test_culled_frame_is_the_dense_frame(ctx, mode, w, h, ns, nc, seed, k):
    ctx.set_scene(*hs.random_scene(max(w, 64), max(h, 64), ns, nc, seed).as_tuple())
    cam = capi.REFERENCE_CAMERA if k == 0 else hs.orbit_camera(w, h, k)
    dense, culled = _both(ctx, w, h, cam, mode)
    for key in ("rgba8", "prim_id", "rgba_i32"):
        assert np.array_equal(dense[key], culled[key]), key
    assert (dense["prim_id"] >= 0).any()


@pytest.mark.parametrize("mode", MODES)
def test_empty_scene_and_single_outputs(ctx, mode):
    ctx.set_scene(*hs.random_scene(640, 480, 0, 0, 1).as_tuple())
    dense, culled = _both(ctx, 203, 77, capi.REFERENCE_CAMERA, mode)
    for key in ("rgba8", "prim_id", "rgba_i32"):
        assert np.array_equal(dense[key], culled[key]), key
    assert (culled["prim_id"] == -1).all()
    ctx.set_scene(*hs.random_scene(640, 480, 30, 30, 3).as_tuple())
    for want in (("rgba8",), ("prim_id",), ("rgba_i32",)):
        dense, culled = _both(ctx, 640, 480, capi.REFERENCE_CAMERA, mode, want=want)
        assert np.array_equal(dense[want[0]], culled[want[0]])


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("which", [1, 2, 3])
def test_reference_scenes(ctx, which, mode):
    ctx.set_scene(*hs.reference_scene(which).as_tuple())
    dense, culled = _both(ctx, 640, 480, capi.REFERENCE_CAMERA, mode)
    for key in ("rgba8", "prim_id", "rgba_i32"):
        assert np.array_equal(dense[key], culled[key]), key


@pytest.mark.parametrize("mode", MODES)
def test_bands_and_device_entry_points(ctx, mode):
    import torch
    w, h = 1920, 1080
    ctx.set_scene(*hs.random_scene(w, h, 64, 64, 2).as_tuple())
    ctx.set_option(capi.RTB_OPT_TILE_CULL, 0)
    full = ctx.render(w, h, want=("rgba8",))["rgba8"]
    ctx.set_option(capi.RTB_OPT_TILE_CULL, mode)
    for world in (2, 3):
        for rank in range(world):
            rows = ctx.band_local_rows(h, 16, rank, world)
            buf = torch.zeros((rows, w, 4), dtype=torch.uint8, device="cuda")
            ctx.render_bands_device(w, h, capi.REFERENCE_CAMERA, 16, rank, world, d_rgba8=buf.data_ptr())
            got = buf.cpu().numpy()
            sel = [y for y in range(h) if (y // 16) % world == rank]
            assert np.array_equal(got, full[sel])
    # whole host frame through the in-launch progress chunks
    frame = np.zeros((h, w, 4), np.uint8)
    ctx.render_bands_host(w, h, capi.REFERENCE_CAMERA, 16, 0, 1, frame)
    assert np.array_equal(frame, full)


def test_culled_kernel_is_much_faster_at_c3(ctx):
    w, h = 3840, 2160
    ctx.set_scene(*hs.random_scene(w, h, 512, 512, 3).as_tuple())
    t = {}
    for mode in (0, 1, 2):
        ctx.set_option(capi.RTB_OPT_TILE_CULL, mode)
        ctx.render(w, h, want=("rgba8",))
        t[mode] = min(ctx.render(w, h, want=("rgba8",))["kernel_ms"] for _ in range(3))
    assert t[1] < t[0] / 4 and t[2] < t[0] / 4, t


def test_unknown_option_is_rejected(ctx):
    with pytest.raises(capi.RtbError) as e:
        ctx.set_option(99, 1)
    assert e.value.status == 1
    with pytest.raises(capi.RtbError):
        ctx.set_option(capi.RTB_OPT_TILE_CULL, 4)


def test_default_mode_is_auto(built_libs):
    """A fresh context renders in the fastest byte-identical mode without being asked (RTB_OPT_TILE_CULL = 3)."""
    with capi.Context(0) as c:
        assert c.get_option(capi.RTB_OPT_TILE_CULL) == capi.TILE_CULL_AUTO
        assert c.get_option(capi.RTB_OPT_DEFER_SCENE_CHECK) == 0
        c.set_scene(*hs.random_scene(640, 480, 100, 100, 7).as_tuple())
        n0 = c.launch_count()
        auto = c.render(640, 480, want=("rgba8",))
        assert c.launch_count() - n0 == 3          # pack + bin_tiles_kernel + raycast_binned_kernel
        c.set_option(capi.RTB_OPT_TILE_CULL, capi.TILE_CULL_DENSE)
        n0 = c.launch_count()
        dense = c.render(640, 480, want=("rgba8",))
        assert c.launch_count() - n0 == 2          # row table + raycast_dense_kernel (the records are already packed)
        assert np.array_equal(auto["rgba8"], dense["rgba8"])
        c.set_option(capi.RTB_OPT_TILE_CULL, capi.TILE_CULL_AUTO)
        c.set_scene(*hs.random_scene(640, 480, 10, 10, 7).as_tuple())   # fewer than 64 primitives: dense
        n0 = c.launch_count()
        c.render(640, 480, want=("rgba8",))
        assert c.launch_count() - n0 == 3          # pack + row table + raycast_dense_kernel


def test_cli_frames_are_identical_in_every_mode(tmp_path):
    """raycast_cli --tile-cull 0|1|2 (RenderState::setTileCull): the PNG files are the same bytes."""
    import os
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = os.path.join(root, "opencl_ray_tracer_b200", "raycast_cli")
    data = []
    for mode in (0, 1, 2, 3):
        out = str(tmp_path / ("m%d.png" % mode))
        r = subprocess.run([exe, "--scene", "random", "--spheres", "80", "--cubes", "70", "--seed", "5", "--width", "800",
                            "--height", "600", "--tile-cull", str(mode), "--out", out], capture_output=True, text=True, timeout=120)
        assert r.returncode == 0, r.stderr
        data.append(open(out, "rb").read())
    assert data[0] == data[1] == data[2] == data[3] and len(data[0]) > 1000
    r = subprocess.run([exe, "--tile-cull", "7"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 2


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("world,band_rows", [(2, 16), (3, 16), (8, 16), (3, 4), (2, 48)])
def test_host_frame_assembled_from_every_ranks_bands(ctx, mode, world, band_rows):
    """rtb_render_bands_host for every rank of a world (one GPU plays them all): the bands, delivered through the
    in-launch progress chunks, make up the dense frame -- tiles that straddle bands and chunk bounds included."""
    w, h = 1000, 700
    ctx.set_scene(*hs.random_scene(w, h, 70, 70, 41).as_tuple())
    cam = hs.orbit_camera(w, h, 7)
    ctx.set_option(capi.RTB_OPT_TILE_CULL, 0)
    full = ctx.render(w, h, cam, want=("rgba8",))["rgba8"]
    ctx.set_option(capi.RTB_OPT_TILE_CULL, mode)
    frame = np.zeros((h, w, 4), np.uint8)
    for rank in range(world):
        ctx.render_bands_host(w, h, cam, band_rows, rank, world, frame)
    assert np.array_equal(frame, full)
