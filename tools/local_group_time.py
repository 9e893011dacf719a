"""rtb_init(n_devices = N) from ONE process (what a C++ MainState would call): end-to-end time of rtb_set_scene + rtb_render
per C3 frame into a pinned host frame, default options and dense, and frames/s of rtb_render_sequence_device (C5) for
N = 1, 2, 4, 8 devices of this box.  One JSON line per N."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from opencl_ray_tracer_b200 import capi, host_scene as hs  # noqa: E402

w, h, ns, nc, seed = 3840, 2160, 512, 512, 3
scene = hs.random_scene(w, h, ns, nc, seed)
arrs = tuple(torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).pin_memory().numpy() for a in scene.as_tuple())
frame = torch.empty((h, w, 4), dtype=torch.uint8).pin_memory().numpy()
ref = None
for n in (1, 2, 4, 8):
    if n > torch.cuda.device_count():
        break
    ctx = capi.Context(0) if n == 1 else capi.Context(devices=list(range(n)))
    ctx.set_option(capi.RTB_OPT_DEFER_SCENE_CHECK, 1)
    out = {"n_devices": n, "frame": "%dx%d, %d prims" % (w, h, ns + nc)}
    for mode, key in ((capi.TILE_CULL_AUTO, "e2e_default_ms"), (capi.TILE_CULL_DENSE, "e2e_dense_ms")):
        ctx.set_option(capi.RTB_OPT_TILE_CULL, mode)
        for _ in range(4):
            ctx.set_scene(*arrs)
            ctx.render(w, h, want=("rgba8",), out={"rgba8": frame})
        t0 = time.perf_counter()
        for _ in range(20):
            ctx.set_scene(*arrs)
            ctx.render(w, h, want=("rgba8",), out={"rgba8": frame})
        out[key] = round((time.perf_counter() - t0) * 1e3 / 20, 4)
        if ref is None:
            ref = frame.copy()
        out[key.replace("_ms", "_frame_matches_1_gpu")] = bool(np.array_equal(frame, ref))
    cams = np.stack([np.concatenate(hs.orbit_camera(w, h, k)) for k in range(120)])
    seq = torch.empty((120, h, w, 4), dtype=torch.uint8, device="cuda:0")
    for mode, key in ((capi.TILE_CULL_DENSE, "c5_dense_frames_per_s"), (capi.TILE_CULL_BINNED, "c5_binned_frames_per_s")):
        ctx.set_option(capi.RTB_OPT_TILE_CULL, mode)
        ctx.render_sequence_device(w, h, cams, seq.data_ptr())
        t0 = time.perf_counter()
        for _ in range(3):
            ctx.render_sequence_device(w, h, cams, seq.data_ptr())
        out[key] = round(3 * 120 / (time.perf_counter() - t0), 1)
    print(json.dumps(out), flush=True)
    del seq
    ctx.close()
