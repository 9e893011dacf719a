"""Where the end-to-end time of one sharded frame goes, per rank (run under torchrun, one process per GPU):
rtb_set_scene (host wall), rtb_group_render_host (host wall; on rank 0 it includes the wait for the other ranks), this
rank's kernel time, and the frame time seen by rank 0.  One JSON line per rank.
  python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/e2e_breakdown_group.py [c3|c4] [dense|binned]"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
from opencl_ray_tracer_b200 import capi, host_scene as hs  # noqa: E402

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
w, h, ns, nc, seed = (7680, 4320, 2048, 2048, 4) if "c4" in sys.argv[1:] else (3840, 2160, 512, 512, 3)
ctx = capi.Context(local)
ctx.set_option(capi.RTB_OPT_TILE_CULL, 2 if "binned" in sys.argv[1:] else 0)
ctx.set_option(capi.RTB_OPT_DEFER_SCENE_CHECK, 1)
scene = hs.random_scene(w, h, ns, nc, seed)
arrs = tuple(torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).pin_memory().numpy() for a in scene.as_tuple())
names = ["rtb_bd_%d_%d" % (os.getpid(), int(time.time() * 1e3) % 100000000)]
dist.broadcast_object_list(names, 0)
grp = capi.Group(ctx, names[0], rank, world, w, h, 16, 3)
ts, tr, tk, tf = [], [], [], []
for i in range(30):
    dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    ctx.set_scene(*arrs)
    t1 = time.perf_counter()
    _, kms = grp.render_host(capi.REFERENCE_CAMERA)
    t2 = time.perf_counter()
    if i >= 8:
        ts.append((t1 - t0) * 1e3); tr.append((t2 - t1) * 1e3); tk.append(kms); tf.append((t2 - t0) * 1e3)
med = lambda a: round(float(np.median(a)), 4)
# back to back, no barrier between frames (what bench.py's e2e loop does)
dist.barrier()
t0 = time.perf_counter()
for i in range(20):
    ctx.set_scene(*arrs)
    grp.render_host(capi.REFERENCE_CAMERA)
dist.barrier()
loop_ms = (time.perf_counter() - t0) * 1e3 / 20
out = {"rank": rank, "world": world, "frame": "%dx%d, %d prims" % (w, h, ns + nc), "tile_cull": ctx.get_option(capi.RTB_OPT_TILE_CULL),
       "rows": ctx.band_local_rows(h, 16, rank, world), "set_scene_ms": med(ts), "render_host_ms": med(tr), "kernel_ms": med(tk),
       "frame_ms_from_barrier": med(tf), "back_to_back_ms_per_frame": round(loop_ms, 4)}
gathered = [None] * world
dist.all_gather_object(gathered, out)
if rank == 0:
    for g in gathered:
        print(json.dumps(g), flush=True)
grp.close()
ctx.close()
dist.destroy_process_group()
