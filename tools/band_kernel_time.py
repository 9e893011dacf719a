"""Kernel time of ONE rank's share of the C3 frame for world = 1, 2, 4, 8 (run on one GPU):
what each GPU of an N-GPU job executes per frame, without the collective."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from opencl_ray_tracer_b200 import capi, host_scene as hs  # noqa: E402

w, h, ns, nc, seed = 3840, 2160, 512, 512, 3
ctx = capi.Context(0)
ctx.set_scene(*hs.random_scene(w, h, ns, nc, seed).as_tuple())
buf = torch.empty((h, w, 4), dtype=torch.uint8, device="cuda")
stream = torch.cuda.Stream()
info = ctx.device_info()
for world in (1, 2, 4, 8):
    rows = ctx.band_local_rows(h, 16, 0, world)
    with torch.cuda.stream(stream):
        for _ in range(5):
            ctx.render_bands_device(w, h, capi.REFERENCE_CAMERA, 16, 0, world, d_rgba8=buf.data_ptr(),
                                    stream=stream.cuda_stream)
        stream.synchronize()
        ctx.kernel_ms_avg(reset=True)
        for _ in range(20):
            ctx.render_bands_device(w, h, capi.REFERENCE_CAMERA, 16, 0, world, d_rgba8=buf.data_ptr(),
                                    stream=stream.cuda_stream)
        stream.synchronize()
    ms, n = ctx.kernel_ms_avg(reset=True)
    print(json.dumps({"world": world, "rows": rows, "kernel_ms": round(ms, 4), "launches": n,
                      "shape": capi.debug_grid_shape(w, rows, info["sm_count"]),
                      "ideal_ms_from_world1": None}))
ctx.close()
