"""rtb_render_device_noalloc (include/rtb.h): one frame with the caller's device buffers in and out and the kernels'
scratch memory (boxes, records, tile words) in a file-scope static __device__ array -- no cudaMalloc in the call.
Its frames must equal, byte for byte (the bar of every integer output here), those of the existing device-pointer entry
point rtb_render_bands_device, on calls in a row that share the scratch array.  No counterpart in the reference.
This file is named to be collected last."""
import numpy as np
import pytest

from opencl_ray_tracer_b200 import capi, host_scene as hs


def _device_scene(torch, scene):
    sph_o, sph_r, sph_c, cube_v, cube_c = scene.as_tuple()
    up = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32).reshape(-1)).cuda()
    return dict(ns=len(np.asarray(sph_r).reshape(-1)), nc=len(np.asarray(cube_c).reshape(-1, 4)),
                sph_o=up(sph_o), sph_r=up(sph_r), sph_c=up(sph_c), cube_v=up(cube_v), cube_c=up(cube_c))


def _ptr(t):
    return t.data_ptr() if t.numel() else 0


def _noalloc(torch, dsc, w, h, cam, out, stream=0, bad=None):
    capi.render_device_noalloc(w, h, cam, dsc["ns"], _ptr(dsc["sph_o"]), _ptr(dsc["sph_r"]), _ptr(dsc["sph_c"]), dsc["nc"],
                               _ptr(dsc["cube_v"]), _ptr(dsc["cube_c"]), out["rgba8"].data_ptr(), out["prim_id"].data_ptr(),
                               out["rgba_i32"].data_ptr(), bad.data_ptr() if bad is not None else 0, stream)


def _outputs(torch, w, h, fill):
    return dict(rgba8=torch.full((h, w), fill, dtype=torch.int32, device="cuda"),
                prim_id=torch.full((h, w), fill, dtype=torch.int32, device="cuda"),
                rgba_i32=torch.full((h, w, 4), fill, dtype=torch.int32, device="cuda"))


def _existing(torch, ctx, scene, w, h, cam):
    """The same frame through the existing entry point (context, cudaMalloc'ed scratch), dense kernel."""
    ctx.set_option(capi.RTB_OPT_TILE_CULL, capi.TILE_CULL_DENSE)
    try:
        ctx.set_scene(*scene.as_tuple())
        out = _outputs(torch, w, h, 0x55)
        ctx.render_bands_device(w, h, cam, 16, 0, 1, out["rgba8"].data_ptr(), out["prim_id"].data_ptr(),
                                out["rgba_i32"].data_ptr())
        torch.cuda.synchronize()
        return {k: v.cpu().numpy() for k, v in out.items()}
    finally:
        ctx.set_option(capi.RTB_OPT_TILE_CULL, capi.TILE_CULL_AUTO)


CASES = [   # two calls in a row (and back again): the second has fewer primitives and tiles than the first left behind
    (1920, 1080, 300, 300, 5, 0),
    (333, 201, 40, 3, 11, 40),
    (3840, 2160, 512, 512, 3, 0),    # the stated maximum: 1024 primitives, 4080 tiles x 32 words
    (640, 480, 0, 0, 1, 0),          # empty scene behind a full one
    (1920, 1080, 300, 300, 5, 0),
]


@pytest.mark.gpu
def test_noalloc_frames_equal_the_existing_entry_point(gpu_ctx):
    import torch
    expected, scenes = [], []
    for (w, h, ns, nc, seed, k) in CASES:
        scene = hs.random_scene(max(w, 64), max(h, 64), ns, nc, seed)
        cam = capi.REFERENCE_CAMERA if k == 0 else hs.orbit_camera(w, h, k)
        scenes.append((scene, cam))
        expected.append(_existing(torch, gpu_ctx, scene, w, h, cam))
    dscs = [_device_scene(torch, s) for s, _ in scenes]
    outs = [_outputs(torch, w, h, 0x33) for (w, h, *_r) in CASES]
    bad = torch.zeros(1, dtype=torch.int32, device="cuda")
    # first call (on the stream the others will use): loads the kernels -- the driver may allocate for code and for the
    # new stream -- from here on free memory must not shrink
    stream = torch.cuda.Stream()
    _noalloc(torch, dscs[0], CASES[0][0], CASES[0][1], scenes[0][1], outs[0], stream=stream.cuda_stream, bad=bad)
    torch.cuda.synchronize()
    outs[0]["prim_id"].fill_(0x33)
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    for i, (w, h, *_r) in enumerate(CASES):          # back to back on one stream, nothing waits in between
        _noalloc(torch, dscs[i], w, h, scenes[i][1], outs[i], stream=stream.cuda_stream, bad=bad)
    stream.synchronize()
    assert torch.cuda.mem_get_info()[0] >= free0, "the call allocated device memory"
    assert int(bad.item()) == 0x7F7F7F7F
    for i, (w, h, ns, nc, *_r) in enumerate(CASES):
        for key in ("rgba8", "prim_id", "rgba_i32"):
            got = outs[i][key].cpu().numpy()
            assert np.array_equal(got, expected[i][key].reshape(got.shape)), (i, key)
        if ns + nc:
            assert (expected[i]["prim_id"] >= 0).any()
    # synchronous form (stream NULL) twice in a row, one output only
    for i in (1, 0):
        w, h = CASES[i][0], CASES[i][1]
        o = _outputs(torch, w, h, 0x11)
        capi.render_device_noalloc(w, h, scenes[i][1], dscs[i]["ns"], _ptr(dscs[i]["sph_o"]), _ptr(dscs[i]["sph_r"]),
                                   _ptr(dscs[i]["sph_c"]), dscs[i]["nc"], _ptr(dscs[i]["cube_v"]), _ptr(dscs[i]["cube_c"]),
                                   d_rgba8=o["rgba8"].data_ptr())
        assert np.array_equal(o["rgba8"].cpu().numpy(), expected[i]["rgba8"].reshape(h, w))
        assert (o["prim_id"].cpu().numpy() == 0x11).all()


@pytest.mark.gpu
def test_noalloc_reports_a_cube_that_is_not_a_box(gpu_ctx):
    import torch
    scene = hs.random_scene(640, 480, 4, 6, 9)
    dsc = _device_scene(torch, scene)
    dsc["cube_v"][144 * 4 + 5] += 3.0    # cube 4, second vertex: no longer an affine image of Cube.cpp's list
    out = _outputs(torch, 320, 240, 0)
    bad = torch.zeros(1, dtype=torch.int32, device="cuda")
    _noalloc(torch, dsc, 320, 240, capi.REFERENCE_CAMERA, out, bad=bad)
    assert int(bad.item()) == 4


@pytest.mark.gpu
def test_noalloc_runs_under_stream_capture(gpu_ctx):
    """Stream capture forbids cudaMalloc: the call captures, and the replayed graph writes the frame."""
    import torch
    w, h = 1280, 720
    scene = hs.random_scene(w, h, 120, 80, 21)
    expected = _existing(torch, gpu_ctx, scene, w, h, capi.REFERENCE_CAMERA)
    dsc = _device_scene(torch, scene)
    out = _outputs(torch, w, h, 0x22)
    _noalloc(torch, dsc, w, h, capi.REFERENCE_CAMERA, out)     # kernels loaded, scratch address known
    for v in out.values():
        v.fill_(0x22)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        _noalloc(torch, dsc, w, h, capi.REFERENCE_CAMERA, out, stream=torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert (out["prim_id"].cpu().numpy() == 0x22).all()        # captured, not run
    for _ in range(2):
        g.replay()
        torch.cuda.synchronize()
        for key in ("rgba8", "prim_id", "rgba_i32"):
            got = out[key].cpu().numpy()
            assert np.array_equal(got, expected[key].reshape(got.shape)), key
        out["prim_id"].fill_(-7)


def test_noalloc_rejects_sizes_beyond_the_static_scratch(built_libs):
    """The limits are checked before anything touches a device: a clear error, also on a box without a GPU."""
    fake = 0x1000   # never dereferenced: the size checks come first
    with pytest.raises(capi.RtbError) as ei:
        capi.render_device_noalloc(640, 480, capi.REFERENCE_CAMERA, 1000, fake, fake, fake, 25, fake, fake, d_rgba8=fake)
    assert ei.value.status == 1 and "RTB_NOALLOC_MAX_PRIMITIVES" in str(ei.value)
    with pytest.raises(capi.RtbError) as ei:   # 8K frame at 1024 primitives: 16 200 tiles x 32 words
        capi.render_device_noalloc(7680, 4320, capi.REFERENCE_CAMERA, 512, fake, fake, fake, 512, fake, fake, d_rgba8=fake)
    assert ei.value.status == 1 and "RTB_NOALLOC_MAX_BIN_WORDS" in str(ei.value)
    with pytest.raises(capi.RtbError) as ei:
        capi.render_device_noalloc(0, 480, capi.REFERENCE_CAMERA, 1, fake, fake, fake, 0, 0, 0, d_rgba8=fake)
    assert ei.value.status == 1
    with pytest.raises(capi.RtbError) as ei:   # float4 arrays must be 16-byte aligned
        capi.render_device_noalloc(64, 48, capi.REFERENCE_CAMERA, 1, fake + 4, fake, fake, 0, 0, 0, d_rgba8=fake)
    assert ei.value.status == 1 and "aligned" in str(ei.value)
