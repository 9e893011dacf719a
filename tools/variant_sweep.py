"""Kernel time of the grid kernel's experiment variants (RTB_GRID_VARIANT, raycast_kernels.cu
launch_grid) on one GPU, each in its own process, with the frame checked against the production
tiling's frame (the bound never changes a pixel, so every variant must give identical bytes).

  python tools/variant_sweep.py [--workload c3] [--variants 0,10,11] [--flush]
"""
import argparse
import hashlib
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

WORKLOADS = {"c1r": (640, 480, 100, 100, 7), "c2": (1920, 1080, 64, 64, 2), "c3": (3840, 2160, 512, 512, 3), "c4": (7680, 4320, 2048, 2048, 4)}


def child(workload, flush, reps):
    import torch
    from opencl_ray_tracer_b200 import capi, host_scene as hs
    w, h, ns, nc, seed = WORKLOADS[workload]
    ctx = capi.Context(0)
    cull = int(os.environ.get("RTB_SWEEP_TILE_CULL", "0"))
    if cull:
        ctx.set_option(capi.RTB_OPT_TILE_CULL, cull)
    ctx.set_scene(*hs.random_scene(w, h, ns, nc, seed).as_tuple())
    buf = torch.zeros((h, w, 4), dtype=torch.uint8, device="cuda")
    fl = torch.empty(256 << 20, dtype=torch.uint8, device="cuda") if flush else None
    stream = torch.cuda.Stream()
    with torch.cuda.stream(stream):
        for _ in range(5):
            ctx.render_bands_device(w, h, capi.REFERENCE_CAMERA, 16, 0, 1, d_rgba8=buf.data_ptr(), stream=stream.cuda_stream)
        stream.synchronize()
        ctx.kernel_ms_avg(reset=True)
        for _ in range(reps):
            if fl is not None:
                fl.zero_()
            ctx.render_bands_device(w, h, capi.REFERENCE_CAMERA, 16, 0, 1, d_rgba8=buf.data_ptr(), stream=stream.cuda_stream)
        stream.synchronize()
    ms, n = ctx.kernel_ms_avg(reset=True)
    digest = hashlib.sha1(buf.cpu().numpy().tobytes()).hexdigest()[:16]
    print(json.dumps({"variant": ({1: "cull", 2: "binned"}[cull] + (":" + os.environ["RTB_BINNED_VARIANT"] if cull == 2 and "RTB_BINNED_VARIANT" in os.environ else "")) if cull else int(os.environ.get("RTB_GRID_VARIANT", "0")), "workload": workload,
                      "kernel_ms": round(ms, 4), "launches": n, "mrays_s": round(w * h / ms / 1e3, 1),
                      "flush": bool(flush), "frame_sha1": digest}))
    ctx.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="c3")
    ap.add_argument("--variants", default="0,1,10,15,20,31")
    ap.add_argument("--flush", action="store_true")
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--child", action="store_true")
    a = ap.parse_args()
    if a.child:
        child(a.workload, a.flush, a.reps)
        return
    ref = None
    for v in a.variants.split(","):
        env = dict(os.environ, RTB_GRID_VARIANT=v)
        if v == "cull" or v.startswith("binned"):  # the tile-culled kernels (RTB_OPT_TILE_CULL = 1 | 2); binned:N = tiling N
            env = dict(os.environ, RTB_GRID_VARIANT="0", RTB_SWEEP_TILE_CULL="1" if v == "cull" else "2")
            if ":" in v:
                env["RTB_BINNED_VARIANT"] = v.split(":")[1]
        cmd = [sys.executable, os.path.abspath(__file__), "--child", "--workload", a.workload, "--reps", str(a.reps)]
        if a.flush:
            cmd.append("--flush")
        out = subprocess.run(cmd, env=env, capture_output=True, text=True)
        line = out.stdout.strip().splitlines()[-1] if out.stdout.strip() else json.dumps({"variant": v, "error": out.stderr[-300:]})
        try:
            j = json.loads(line)
            if ref is None and "frame_sha1" in j:
                ref = j["frame_sha1"]
            j["same_frame_as_first"] = (j.get("frame_sha1") == ref)
            line = json.dumps(j)
        except Exception:
            pass
        print(line, flush=True)


if __name__ == "__main__":
    main()
