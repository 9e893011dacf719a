"""Kernel time of the dense kernel (RTB_OPT_TILE_CULL = 0) for ONE rank's share of a frame at world = 1, 2, 4, 8
(what each GPU of an N-GPU job executes per frame), L2 flushed between launches, for the tilings behind
RTB_DENSE_TILING (-1 = the cost model's pick) and -- with librtb_exp.so (RTB_BUILD_EXP=1 build.sh) -- the round-1
kernel behind RTB_GRID_VARIANT.  One process per configuration; every frame is hashed so that tilings can be compared.
  python tools/dense_sweep.py [--workloads c3,c4,c2] [--tilings 0,1,2,3]  (experiments build: 0 = the product's 2-warp CTA, 1 = 4-warp, 2 = split 2, 3 = split 4) [--old 10,20]"""
import argparse
import hashlib
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
WORKLOADS = {"c2": (1920, 1080, 64, 64, 2), "c3": (3840, 2160, 512, 512, 3), "c4": (7680, 4320, 2048, 2048, 4)}


def child(workloads, worlds):
    import torch
    from opencl_ray_tracer_b200 import capi, host_scene as hs
    ctx = capi.Context(0)
    ctx.set_option(capi.RTB_OPT_TILE_CULL, 0)
    info = ctx.device_info()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    stream = torch.cuda.Stream()
    for wl in workloads:
        w, h, ns, nc, seed = WORKLOADS[wl]
        ctx.set_scene(*hs.random_scene(w, h, ns, nc, seed).as_tuple())
        buf = torch.zeros((h, w, 4), dtype=torch.uint8, device="cuda")
        for world in worlds:
            rows = ctx.band_local_rows(h, 16, 0, world)
            n = 20 if wl != "c4" else 8
            with torch.cuda.stream(stream):
                for _ in range(3):
                    ctx.render_bands_device(w, h, capi.REFERENCE_CAMERA, 16, 0, world, d_rgba8=buf.data_ptr(), stream=stream.cuda_stream)
                stream.synchronize()
                ctx.kernel_ms_avg(reset=True)
                for _ in range(n):
                    flush.zero_()
                    ctx.render_bands_device(w, h, capi.REFERENCE_CAMERA, 16, 0, world, d_rgba8=buf.data_ptr(), stream=stream.cuda_stream)
                stream.synchronize()
            ms, k = ctx.kernel_ms_avg(reset=True)
            digest = hashlib.sha1(buf[:rows].cpu().numpy().tobytes()).hexdigest()[:12]
            tests = rows * w * (ns + nc)
            print(json.dumps({"lib": os.path.basename(os.environ.get("RTB_LIB", "librtb.so")), "tiling": os.environ.get("RTB_DENSE_TILING", "-1"),
                              "old_variant": os.environ.get("RTB_GRID_VARIANT", ""), "workload": wl, "world": world, "rows": rows,
                              "kernel_ms": round(ms, 4), "launches": k, "tests_per_s": round(tests / (ms * 1e-3), 0) if ms > 0 else None,
                              "dispatch_frac": round(tests / (ms * 1e-3) / (info["sm_count"] * 4 * 32 * 1.965e9 / 2), 4) if ms > 0 else None,
                              "shape": capi.debug_grid_shape(w, rows, info["sm_count"]), "sha1": digest}), flush=True)
        del buf
    ctx.close()


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--workloads", default="c3")
    ap.add_argument("--worlds", default="1,2,4,8")
    ap.add_argument("--tilings", default="-1")
    ap.add_argument("--old", default="")
    ap.add_argument("--libs", default="", help="comma list of librtb_<name>.so builds (compile-time variants) to time as well")
    ap.add_argument("--child", action="store_true")
    a = ap.parse_args()
    if a.child:
        child(a.workloads.split(","), [int(x) for x in a.worlds.split(",")])
    else:
        exp = os.path.join(ROOT, "opencl_ray_tracer_b200", "librtb_exp.so")
        # the product has ONE tiling (0); the others are compiled into the experiments build only
        runs = [dict(RTB_DENSE_TILING=t, **({} if t in ("-1", "0") else {"RTB_LIB": exp})) for t in a.tilings.split(",") if t]
        runs += [dict(RTB_GRID_VARIANT=v, RTB_LIB=exp) for v in a.old.split(",") if v]
        runs += [dict(RTB_LIB=os.path.join(ROOT, "opencl_ray_tracer_b200", "librtb_%s.so" % n)) for n in a.libs.split(",") if n]
        for env in runs:
            out = subprocess.run([sys.executable, os.path.abspath(__file__), "--child", "--workloads", a.workloads, "--worlds", a.worlds],
                                 env=dict(os.environ, **env), capture_output=True, text=True)
            print(out.stdout.strip() if out.stdout.strip() else out.stderr[-600:], flush=True)
