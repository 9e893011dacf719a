"""Where the end-to-end time of one C3 frame goes: rtb_set_scene, rtb_render (host wall), the device
span rtb_render reports (total_ms: pack + kernel + copies) and the grid kernel alone."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from opencl_ray_tracer_b200 import capi, host_scene as hs  # noqa: E402

w, h, ns, nc, seed = 3840, 2160, 512, 512, 3
if "c4" in sys.argv[1:]:
    w, h, ns, nc, seed = 7680, 4320, 2048, 2048, 4
ctx = capi.Context(0)
if "defer" in sys.argv[1:]:
    ctx.set_option(capi.RTB_OPT_DEFER_SCENE_CHECK, 1)
if "dense" in sys.argv[1:]:
    ctx.set_option(capi.RTB_OPT_TILE_CULL, 0)
elif "binned" in sys.argv[1:]:
    ctx.set_option(capi.RTB_OPT_TILE_CULL, 2)
elif "cull" in sys.argv[1:]:
    ctx.set_option(capi.RTB_OPT_TILE_CULL, 1)
scene = hs.random_scene(w, h, ns, nc, seed)
arrs = scene.as_tuple()
buf, ptr = capi.host_alloc(w * h * 4)
frame = buf.reshape(h, w, 4)
ts, tr, tot, ker = [], [], [], []
for i in range(25):
    t0 = time.perf_counter()
    ctx.set_scene(*arrs)
    t1 = time.perf_counter()
    r = ctx.render(w, h, want=("rgba8",), out={"rgba8": frame})
    t2 = time.perf_counter()
    if i >= 5:
        ts.append((t1 - t0) * 1e3); tr.append((t2 - t1) * 1e3); tot.append(r["total_ms"]); ker.append(r["kernel_ms"])
med = lambda a: float(np.median(a))
print(json.dumps({"frame": "%dx%d, %d prims" % (w, h, ns + nc), "tile_cull": ctx.get_option(capi.RTB_OPT_TILE_CULL), "set_scene_ms": round(med(ts), 4), "render_wall_ms": round(med(tr), 4),
                  "render_device_span_ms": round(med(tot), 4), "grid_kernel_ms": round(med(ker), 4),
                  "e2e_ms": round(med(ts) + med(tr), 4)}))
# same with the scene left resident (render only)
tr2 = []
for i in range(15):
    t1 = time.perf_counter()
    r = ctx.render(w, h, want=("rgba8",), out={"rgba8": frame})
    tr2.append((time.perf_counter() - t1) * 1e3)
print(json.dumps({"render_only_wall_ms": round(med(tr2[3:]), 4)}))
capi.host_free(ptr)
ctx.close()
