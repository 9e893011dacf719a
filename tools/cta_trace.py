"""CTA timeline of the dense kernel (experiments build: RTB_BUILD_EXP=1 opencl_ray_tracer_b200/build.sh): where a launch's
time goes besides the dense loop -- start-up ramp, wait for the first row-table chunk, ragged end.
  python tools/cta_trace.py [--workload c3] [--worlds 1,8] [--tilings 0,1]
One JSON line per (tiling, world): launch span, CTA lifetime, first-chunk wait, when the first / last CTA started,
busy fraction of the SMs, tail (time during which fewer than half of the SMs still had a CTA)."""
import argparse
import json
import os
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
WORKLOADS = {"c2": (1920, 1080, 64, 64, 2), "c3": (3840, 2160, 512, 512, 3), "c4": (7680, 4320, 2048, 2048, 4)}


def child(wl, world):
    import torch
    from opencl_ray_tracer_b200 import capi, host_scene as hs
    w, h, ns, nc, seed = WORKLOADS[wl]
    ctx = capi.Context(0)
    ctx.set_option(capi.RTB_OPT_TILE_CULL, 0)
    ctx.set_scene(*hs.random_scene(w, h, ns, nc, seed).as_tuple())
    buf = torch.zeros((h, w, 4), dtype=torch.uint8, device="cuda")
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    for _ in range(4):
        flush.zero_()
        torch.cuda.synchronize()
        ctx.render_bands_device(w, h, capi.REFERENCE_CAMERA, 16, 0, world, d_rgba8=buf.data_ptr())
    ctx.close()


def summarise(path):
    launches, cur = [], None
    for ln in open(path):
        if ln.startswith("launch"):
            cur = []
            launches.append(cur)
        else:
            cur.append([int(x) for x in ln.split()])
    a = np.array(launches[-1], dtype=np.int64)           # the last launch (warm)
    t0 = a[:, 0].min()
    start, first, end, sm = (a[:, 0] - t0) / 1e3, (a[:, 1] - t0) / 1e3, (a[:, 2] - t0) / 1e3, a[:, 3]
    span = end.max()
    life = end - start
    sms = np.unique(sm)
    # number of SMs that still have a CTA at time t
    last_end_per_sm = np.array([end[sm == s].max() for s in sms])
    busy = sum((end[sm == s] - start[sm == s]).sum() for s in sms)
    per_sm_ctas = np.array([(sm == s).sum() for s in sms])
    return {"ctas": int(len(a)), "sms_used": int(len(sms)), "span_us": round(float(span), 2),
            "cta_life_us": {"median": round(float(np.median(life)), 2), "min": round(float(life.min()), 2), "max": round(float(life.max()), 2)},
            "first_chunk_wait_us_median": round(float(np.median(first - start)), 2),
            "last_cta_start_us": round(float(start.max()), 2),
            "start_spread_of_first_wave_us": round(float(np.sort(start)[min(len(start), 592) - 1]), 2),
            "ctas_per_sm": {"min": int(per_sm_ctas.min()), "max": int(per_sm_ctas.max())},
            "sm_finish_us": {"min": round(float(last_end_per_sm.min()), 2), "median": round(float(np.median(last_end_per_sm)), 2),
                             "max": round(float(last_end_per_sm.max()), 2)},
            "tail_us_fewer_than_half_the_sms_busy": round(float(span - np.sort(last_end_per_sm)[len(sms) // 2]), 2)}


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="c3")
    ap.add_argument("--worlds", default="1,8")
    ap.add_argument("--tilings", default="0,1")
    ap.add_argument("--child", default="")
    a = ap.parse_args()
    if a.child:
        wl, world = a.child.split(",")
        child(wl, int(world))
        sys.exit(0)
    exp = os.path.join(ROOT, "opencl_ray_tracer_b200", "librtb_exp.so")
    for t in a.tilings.split(","):
        for world in a.worlds.split(","):
            path = tempfile.mktemp(prefix="cta_trace_")
            env = dict(os.environ, RTB_LIB=exp, RTB_DENSE_TILING=t, RTB_CTA_TRACE=path)
            out = subprocess.run([sys.executable, os.path.abspath(__file__), "--child", "%s,%s" % (a.workload, world)], env=env,
                                 capture_output=True, text=True)
            if not os.path.isfile(path):
                print(json.dumps({"tiling": t, "world": world, "error": out.stderr[-300:]}))
                continue
            print(json.dumps(dict({"workload": a.workload, "tiling": int(t), "world": int(world)}, **summarise(path))), flush=True)
            os.remove(path)
