"""Kernel time of ONE rank's share of the C3 frame for world = 1, 2, 4, 8 (run on one GPU):
what each GPU of an N-GPU job executes per frame, without the fence.  With --variants the
tilings behind RTB_GRID_VARIANT are timed too (one process each) and every frame is checked
against the production tiling's."""
import argparse
import hashlib
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def child():
    import torch
    from opencl_ray_tracer_b200 import capi, host_scene as hs
    w, h, ns, nc, seed = 3840, 2160, 512, 512, 3
    ctx = capi.Context(0)
    ctx.set_scene(*hs.random_scene(w, h, ns, nc, seed).as_tuple())
    buf = torch.zeros((h, w, 4), dtype=torch.uint8, device="cuda")
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
        digest = hashlib.sha1(buf[:rows].cpu().numpy().tobytes()).hexdigest()[:12]
        print(json.dumps({"variant": int(os.environ.get("RTB_GRID_VARIANT", "0")), "world": world, "rows": rows,
                          "kernel_ms": round(ms, 4), "launches": n,
                          "production_shape": capi.debug_grid_shape(w, rows, info["sm_count"]), "sha1": digest}), flush=True)
    ctx.close()


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--variants", default="0")
    ap.add_argument("--child", action="store_true")
    a = ap.parse_args()
    if a.child:
        child()
    else:
        for v in a.variants.split(","):
            out = subprocess.run([sys.executable, os.path.abspath(__file__), "--child"],
                                 env=dict(os.environ, RTB_GRID_VARIANT=v), capture_output=True, text=True)
            print(out.stdout.strip() if out.stdout.strip() else out.stderr[-400:], flush=True)
