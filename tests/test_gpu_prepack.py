"""rtb_prepack_cameras / the pipelined rtb_render_sequence_device: records packed ahead of their frames on the
context's side stream must be the records a frame's own pack would have produced -- same bytes out -- and must
never outlive the scene or leak into a frame of another size."""
import numpy as np
import pytest

from opencl_ray_tracer_b200 import capi, host_scene as hs

pytestmark = pytest.mark.gpu

W, H = 640, 360


def _flat(cams):
    return np.stack([np.concatenate(c) for c in cams]).astype(np.float32)


@pytest.mark.parametrize("mode", [0, 2])
def test_prepacked_cameras_give_the_same_frames(gpu_ctx, mode):
    ctx = gpu_ctx
    ctx.set_option(capi.RTB_OPT_TILE_CULL, mode)
    try:
        ctx.set_scene(*hs.random_scene(W, H, 60, 50, 21).as_tuple())
        cams = [hs.orbit_camera(W, H, k) for k in (0, 5, 17, 44, 90)]
        direct = [ctx.render(W, H, c, want=("rgba8", "prim_id")) for c in cams]
        ctx.prepack_cameras(W, H, _flat(cams))
        per_frame = 2 if mode == 2 else 1          # launches of one render without its pack
        for c, d in reversed(list(zip(cams, direct))):
            n0 = ctx.launch_count()
            again = ctx.render(W, H, c, want=("rgba8", "prim_id"))
            assert ctx.launch_count() - n0 == per_frame, "a prepacked camera must not be packed again"
            assert np.array_equal(again["rgba8"], d["rgba8"]) and np.array_equal(again["prim_id"], d["prim_id"])
        # a camera outside the set, and another frame size, pack for themselves
        other = hs.orbit_camera(W, H, 33)
        n0 = ctx.launch_count()
        ctx.render(W, H, other, want=("rgba8",))
        assert ctx.launch_count() - n0 == per_frame + (1 if mode == 2 else 2)   # + pack (+ row table for the dense kernel)
        small = ctx.render(320, 200, cams[1], want=("rgba8",))["rgba8"]
        ctx.prepack_cameras(W, H, np.zeros((0, 16), np.float32))   # drops the slots
        assert np.array_equal(ctx.render(320, 200, cams[1], want=("rgba8",))["rgba8"], small)
    finally:
        ctx.set_option(capi.RTB_OPT_TILE_CULL, capi.TILE_CULL_AUTO)


def test_new_scene_drops_the_slots(gpu_ctx):
    ctx = gpu_ctx
    cams = [hs.orbit_camera(W, H, k) for k in (3, 9)]
    ctx.set_scene(*hs.random_scene(W, H, 40, 40, 1).as_tuple())
    ctx.prepack_cameras(W, H, _flat(cams))
    a = ctx.render(W, H, cams[0], want=("rgba8",))["rgba8"]
    ctx.set_scene(*hs.random_scene(W, H, 40, 40, 2).as_tuple())
    b = ctx.render(W, H, cams[0], want=("rgba8",))["rgba8"]
    assert not np.array_equal(a, b)
    ctx.set_scene(*hs.random_scene(W, H, 40, 40, 1).as_tuple())
    assert np.array_equal(ctx.render(W, H, cams[0], want=("rgba8",))["rgba8"], a)


@pytest.mark.parametrize("mode", [0, 2])
def test_sequence_equals_frame_by_frame(gpu_ctx, mode):
    import torch
    ctx = gpu_ctx
    ctx.set_option(capi.RTB_OPT_TILE_CULL, mode)
    try:
        ctx.set_scene(*hs.random_scene(W, H, 60, 50, 23).as_tuple())
        cams = [hs.orbit_camera(W, H, k) for k in range(0, 120, 9)]
        out = torch.zeros((len(cams), H, W, 4), dtype=torch.uint8, device="cuda")
        ctx.render_sequence_device(W, H, _flat(cams), out.data_ptr())
        got = out.cpu().numpy()
        for i, c in enumerate(cams):
            assert np.array_equal(got[i], ctx.render(W, H, c, want=("rgba8",))["rgba8"]), i
        with pytest.raises(capi.RtbError):
            bad = _flat(cams)
            bad[2, 4:8] = 0.0     # dU = 0: degenerate camera
            ctx.render_sequence_device(W, H, bad, out.data_ptr())
        assert np.array_equal(ctx.render(W, H, cams[1], want=("rgba8",))["rgba8"], got[1])   # context still usable
    finally:
        ctx.set_option(capi.RTB_OPT_TILE_CULL, capi.TILE_CULL_AUTO)


@pytest.mark.parametrize("mode", [0, 2])
def test_long_sequence_on_two_lanes(gpu_ctx, mode):
    """rtb_render_sequence_device: more frames than camera slots (256), consecutive frames on alternating streams with
    their own bin words; every 23rd frame (and the last) must equal the single-frame render of its camera."""
    import torch
    ctx = gpu_ctx
    w, h, nf = 200, 120, 300
    ctx.set_option(capi.RTB_OPT_TILE_CULL, mode)
    try:
        ctx.set_scene(*hs.random_scene(w, h, 70, 30, 5).as_tuple())
        cams = [hs.orbit_camera(w, h, k % 120) for k in range(nf)]
        out = torch.zeros((nf, h, w, 4), dtype=torch.uint8, device="cuda")
        torch.cuda.synchronize()
        ctx.render_sequence_device(w, h, _flat(cams), out.data_ptr())
        seq = out.cpu().numpy()
        for k in list(range(0, nf, 23)) + [nf - 2, nf - 1]:
            assert np.array_equal(seq[k], ctx.render(w, h, cams[k], want=("rgba8",))["rgba8"]), k
    finally:
        ctx.set_option(capi.RTB_OPT_TILE_CULL, capi.TILE_CULL_AUTO)
