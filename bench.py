#!/usr/bin/env python3
"""bench.py -- headline benchmark of the ray-cast path (BASELINE.json metric: Mrays/s and
ray-primitive tests/s on synthetic scenes; % of FP32 peak; the CPU path timed beside it).

  python bench.py --gpus N --steps K --warmup W            our arm (sm_100a kernels through the C ABI)
  python bench.py --impl reference --gpus N --steps K ...  the reference's CPU path on the host cores

One "step" = one frame of the workload.  With N > 1 (one process per GPU, torchrun) the frame is
sharded as interleaved 16-row bands (strong scaling: the frame is fixed); every rank's kernel stores its
bands straight into rank 0's frame over NVLink peer memory and the launch's last CTA signals rank 0
(--no-peer: bands gathered over NCCL and de-interleaved by one copy kernel).  The headline numbers are
the dense O(rays x primitives) kernel's; a second pass in the same run times the binned tile-cull mode
("binned", byte-identical frames).  Prints ONE JSON line on rank 0.  Nothing here reads /root/reference.
"""
import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

# BASELINE.json configs: name -> (width, height, spheres, cubes, seed)
WORKLOADS = {
    "c1": (640, 480, 2, 4, 1),             # the reference's default scene (createScene1) at its window size; `seed` = scene id
    "c2": (1920, 1080, 64, 64, 2),
    "c3": (3840, 2160, 512, 512, 3),       # the roofline point: 4K x 1024 primitives
    "c4": (7680, 4320, 2048, 2048, 4),
}
SEQUENCE = {"c5": ("c3", 120, 3)}           # BASELINE config 5: 120-frame orbit (3 degrees per frame) of C3's scene
BAND_ROWS = 16
GROUP_SLOTS = 3   # frame slots of a multi-GPU job (rtb_group_join): the ranks run up to 2 frames ahead of rank 0

# FP32 operations the shipped kernel executes per ray-primitive test (FMA = 2; min / compare / bit
# operations / votes count 0).  Cubes and spheres go through the same dense loop: per ray ONE FMA
# (w = sA*u + rowW, issued as FFMA2 on ray pairs) = 2 flop, then |w| is folded into a running minimum m (3-input
# FMNMX per ray pair); per thread and primitive one more FMA, q = m*m + rowC (compared with 2), shared by the
# thread's rays (2 flop / 16 rays at the roofline point).  The row terms (rowW: 1 FFMA; rowC: 1 FADD + 2 FMUL =
# 5 flop) are computed once per primitive and CTA pixel row into the shared-memory row table, i.e. amortised over
# the width of a CTA's pixel tile (256 rays at the roofline point: +0.02 flop per test).  The reference-exact slab /
# sphere arithmetic runs only on rays inside a bound (a few per mille of the tests) and is NOT counted.
# Up to v7 the kernel evaluated q per ray (a second FMA: 4.02 flop per test, 0.704 ms at C3 = 68.5 % of the FP32
# peak).  q is monotone in |w|, so the minimum over a thread's rays is the q of the smallest |w|: the same decision
# bit for bit with half the FMAs, and the frame takes 0.585 ms -- the FP32 fraction of the shipped kernel is LOWER
# while every number a user sees is better.  What bounds it is the dispatch port: an FFMA2 and an FMNMX3 each hold
# it for 2 clocks and serve 2 rays, i.e. 2 clocks per test and scheduler ("dispatch" below).
FLOP_PER_TEST = 2.0              # round 2: w' = fmaf(u, sAn, rWn), ONE FMA per ray-primitive test; nothing else is FP32 arithmetic
PAIR_CLOCKS_FALLBACK = 5.0       # scheduler clocks per (FFMA2 + FMNMX3) pair if the probe binary is missing


def bench_config(workload, world):
    """The `config` object of a bench line: the SAME dict in our arm and in the reference arm (same_config)."""
    return {"workload": workload_name(workload) if workload in WORKLOADS else
            "c5: %d-frame orbit (%d degrees per frame about the scene centre) of %s" % (
                SEQUENCE[workload][1], SEQUENCE[workload][2], workload_name(SEQUENCE[workload][0])),
            "parallelism": "interleaved %d-row bands over %d GPU(s); one process per GPU, one C call per frame and rank "
                           "(rtb_group_render: kernels store into rank 0's frame slots over NVLink peer memory, per-rank done / "
                           "credit words, no collective in the frame loop); NCCL broadcasts the scene" % (BAND_ROWS, world)
                           if world > 1 else "one GPU, whole frame per launch",
            "l2": "256 MiB buffer written between timed iterations (L2 flush); per-step CUDA events on the launching stream",
            "records": "static camera: the per-frame record pack / row table are cached after the first frame "
                       "(moving_camera_ms_per_step times the same loop with a new camera every step)"}


def make_scene(key):
    """The workload's scene (host arrays in the reference's layout)."""
    from opencl_ray_tracer_b200 import host_scene as hs
    w, h, ns, nc, seed = WORKLOADS[key]
    return hs.reference_scene(1) if key == "c1" else hs.random_scene(w, h, ns, nc, seed)


def workload_name(key):
    w, h, ns, nc, seed = WORKLOADS[key]
    if key == "c1":
        return "c1: %dx%d, the reference's default scene (createScene1, MainState.cpp:419-462), reference camera" % (w, h)
    return "%s: %dx%d, %d spheres + %d cubes, createScene3 recipe, seed %d, reference camera" % (key, w, h, ns, nc, seed)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c3", choices=sorted(WORKLOADS) + sorted(SEQUENCE))
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="target CPU time of the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-c4", action="store_true", help="skip the BASELINE config 4 sub-record of a c3 run")
    ap.add_argument("--no-binned", action="store_true", help="skip the second pass with the binned tile cull (RTB_OPT_TILE_CULL = 2)")
    ap.add_argument("--lanes", type=int, default=2, help="c5 at N > 1: contexts (+ groups, streams) per rank that the frames alternate between")
    return ap.parse_args()


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.isfile(p):
        return json.load(open(p)), "measured (MEASURED_PEAKS.json)"
    return {"hbm_gbs": 6650.0, "sm_max_mhz": 1965.0}, "fallback (B200_PROFILING.md)"


class ClockSampler(threading.Thread):
    """Samples SM clock + throttle reasons through NVML while the timed region runs."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.stop_flag = index, [], set(), False
        self.max_mhz = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {
            nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
            nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
            nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
            nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap",
            nv.nvmlClocksThrottleReasonHwPowerBrakeSlowdown: "hw_power_brake",
        }
        while not self.stop_flag:
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.005)

    def result(self):
        s = sorted(self.samples)
        return {"sm_mhz": (s[len(s) // 2] if s else None), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(s)}


def ncu_traffic_bytes(workload, key=None):
    """dram__bytes_read.sum + dram__bytes_write.sum of one launch of the dominant kernel (key = "binned": of the
    two launches of a binned frame), from the committed ncu --set full captures (profiles/r02_ncu_traffic.json);
    None for other workloads."""
    try:
        j = json.load(open(os.path.join(ROOT, "profiles", "r02_ncu_traffic.json")))
        if j.get("workload") == workload:
            if key:
                j = j[key]
            return j["dram_bytes_read"] + j["dram_bytes_write"]
    except Exception:
        pass
    return None


def fp32_probe_tflops():
    """Empirical FP32 peak: profiles/microbench/fp32_issue_probe (packed-FFMA2 chains), TFLOP/s."""
    import subprocess
    exe = os.path.join(ROOT, "profiles", "microbench", "fp32_issue_probe")
    if not os.path.isfile(exe):
        return None
    try:
        out = subprocess.run([exe], capture_output=True, text=True, timeout=60).stdout
        sms = 148
        for line in out.splitlines():
            j = json.loads(line)
            if "sms" in j:
                sms = j["sms"]
            if j.get("mix") == "8xFFMA2":
                fmas = 16.0 * 4096 * 1024 * sms  # per launch: 16 FMA x ITERS x threads/SM x SMs
                return 2.0 * fmas / (j["ms"] * 1e-3) / 1e12
    except Exception:
        return None
    return None


def cpu_baseline_port(scene, w, h, target_seconds):
    """The oracle port (hoisted copies, OpenMP over rows, bit-identical to the reference's CPU path)
    on all host cores, on a bounded sample of rows of the SAME frame."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from oracle.oracle import PortOracle, Scene
    port = PortOracle()
    sc = Scene(*scene.as_tuple())
    cores = port.max_threads()
    probe_rows = max(cores, 8)
    step = max(1, h // probe_rows)
    t0 = time.perf_counter()
    port.render(sc, w, h, rows=(0, h, step))
    dt = time.perf_counter() - t0
    n_probe = len(range(0, h, step))
    rows = int(min(h, max(n_probe, n_probe * target_seconds / max(dt, 1e-6))))
    step = max(1, h // rows)
    t0 = time.perf_counter()
    port.render(sc, w, h, rows=(0, h, step))
    dt = time.perf_counter() - t0
    n_rows = len(range(0, h, step))
    return {"value": n_rows * w / dt / 1e6, "unit": "Mrays/s", "cores": cores, "kind": "port",
            "sample": "rows 0,%d,... (%d of %d rows) of the same %dx%d frame, %.1f s" % (step, n_rows, h, w, h, dt),
            "tests_per_s": n_rows * w * (scene.ns + scene.nc) / dt}


def cpu_faithful_frame(key, scene, w, h, gpu_rgba8):
    """BASELINE.md s3 `cpu_faithful`: the reference's own ray loop (oracle/_ref: MainState.cpp:941-956 with collide /
    intersectTri / intersectSphere / Cube compiled from its sources; else the port's faithful mode), ONE thread, the WHOLE
    frame, timed beside the GPU line; the GPU frame is compared with it.  c1 also goes to PNG: the CPU frame through the
    reference's lodepng (MainState.cpp:410-417), the GPU frame through the product's writer, both decoded by lodepng."""
    import tempfile
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from parity import default_origins
    from oracle.oracle import PortOracle, RefOracle, Scene, ref_available
    d = np.array([0, 0, -1, -1], np.float32)
    kind = "reference" if (ref_available() and (key == "c1" or scene.cube_recipes is not None)) else "port"
    ref = RefOracle() if ref_available() else None
    if kind == "reference":
        if key == "c1":
            ref.build_scene(1)
        else:
            ref.clear_scene()
            for o, r, c in zip(scene.sph_o, scene.sph_r, scene.sph_c):
                ref.add_sphere(o, r, c)
            for c, q in zip(scene.cube_c, scene.cube_recipes):   # createScene3's loop body, MainState.cpp:618-638
                ref.add_cube_ops(c, [("scale", (q[0],) * 3), ("rotate", (0, 0, q[1])), ("rotate", (0, q[2], 0)),
                                     ("rotate", (q[3], 0, 0)), ("translate", (q[4], q[5], q[6]))])
        o = default_origins(w, h)
        t0 = time.perf_counter()
        rgba = ref.render(o, d, shape=(h, w, 4))
        dt = time.perf_counter() - t0
    else:
        port = PortOracle()
        t0 = time.perf_counter()
        rgba = port.render(Scene(*scene.as_tuple()), w, h, faithful=True, threads=1)[0]
        dt = time.perf_counter() - t0
    cpu8 = (rgba & 0xFF).astype(np.uint8)
    diff = np.abs(cpu8[..., :3].astype(np.int32) - gpu_rgba8[..., :3].astype(np.int32))
    diff = np.minimum(diff, 256 - diff).max(-1)
    out = {"value": w * h / dt / 1e6, "unit": "Mrays/s", "cores": 1, "kind": kind, "seconds": dt,
           "sample": "the whole %dx%d frame, 1 thread (the reference is 'Non-Parallel CPU')" % (w, h),
           "tests_per_s": w * h * (scene.ns + scene.nc) / dt,
           "gpu_frame_vs_this": {"rgb_within1_frac": float((diff <= 1).mean()), "rgb_exact_frac": float((diff == 0).mean()),
                                 "pixels_off_by_more": int((diff > 1).sum()), "bar": "tests/parity.py: >= 0.999 within 1, the rest on silhouettes"}}
    if key == "c1" and ref is not None:
        from opencl_ray_tracer_b200 import host_scene as hs
        tmp = tempfile.mkdtemp(prefix="rtb_c1_png_")
        a, b = os.path.join(tmp, "reference_lodepng.png"), os.path.join(tmp, "product_writer.png")
        t0 = time.perf_counter()
        rc = ref.lib.ref_png_encode(a.encode(), np.ascontiguousarray(cpu8).ctypes.data, w, h)
        t_ref = time.perf_counter() - t0
        t0 = time.perf_counter()
        hs.png_write(b, gpu_rgba8)
        t_own = time.perf_counter() - t0
        out["png"] = {"what": "CPU frame -> lodepng::encode (the reference's call); GPU frame -> the product's encodePNG; both decoded by lodepng",
                      "lodepng_rc": int(rc), "lodepng_bytes": os.path.getsize(a), "lodepng_ms": t_ref * 1e3,
                      "product_bytes": os.path.getsize(b), "product_ms": t_own * 1e3,
                      "lodepng_file_decodes_to_cpu_frame": bool(np.array_equal(ref.png_decode(a), cpu8)),
                      "product_file_decodes_to_gpu_frame": bool(np.array_equal(ref.png_decode(b), gpu_rgba8))}
        for f in (a, b):
            os.remove(f)
        os.rmdir(tmp)
    return out


def _setup_dist():
    import torch
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    if world > 1 and not dist.is_initialized():
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    return rank, world, local_rank


def _broadcast_scene(scene, rank, world):
    """The scene lives on rank 0's host: the flattened arrays go to the other ranks over NCCL (north_star)."""
    import torch
    import torch.distributed as dist
    from opencl_ray_tracer_b200 import host_scene as hs
    if world == 1:
        return scene
    arrays = []
    for a in scene.as_tuple():
        g = torch.from_numpy(np.ascontiguousarray(a) if rank == 0 else np.zeros_like(a)).cuda()
        dist.broadcast(g, 0)
        arrays.append(g.cpu().numpy())
    out = hs.SceneArrays(*arrays, label=scene.label)
    out.cube_recipes = scene.cube_recipes
    return out


def _group_name(tag, rank, world):
    """One name per job and workload for the ranks' shared-memory control block (rank 0 picks it)."""
    import torch.distributed as dist
    names = ["rtb_%s_%d_%d" % (tag, os.getpid(), int(time.time() * 1e3) % 100000000)]
    if world > 1:
        dist.broadcast_object_list(names, 0)
    return names[0]


def measure_workload(args, key, ctx, rank, world, local_rank, steps, with_binned, with_moving):
    """Times one BASELINE workload on the job's GPUs: device-resident frames (`value`), frames delivered to the host
    through the C ABI (`e2e`), for the dense O(rays x primitives) kernel and -- with_binned -- for the binned mode.
    world > 1: ONE C call per frame and rank (rtb_group_render / rtb_group_render_host): no collective, no Python
    orchestration inside the frame."""
    import torch
    import torch.distributed as dist
    from opencl_ray_tracer_b200 import capi, host_scene as hs

    w, h, ns, nc, seed = WORKLOADS[key]
    scene = _broadcast_scene(make_scene(key), rank, world)
    ns, nc = scene.ns, scene.nc
    ctx.set_option(capi.RTB_OPT_DEFER_SCENE_CHECK, 0)
    ctx.set_scene(*scene.as_tuple())
    cam = capi.REFERENCE_CAMERA
    n_px = w * h
    my_rows = ctx.band_local_rows(h, BAND_ROWS, rank, world)
    stream = torch.cuda.Stream()          # everything below (kernels, events, the L2 flush) runs on this stream
    torch.cuda.set_stream(stream)
    sptr = stream.cuda_stream
    grp = capi.Group(ctx, _group_name(key, rank, world), rank, world, w, h, BAND_ROWS, GROUP_SLOTS) if world > 1 else None
    d_frame = torch.empty((h, w, 4), dtype=torch.uint8, device="cuda") if (world == 1 or rank == 0) else None
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")  # > 126 MB L2
    h_frame = torch.empty((h, w, 4), dtype=torch.uint8).pin_memory().numpy() if world == 1 else None
    # the step's inputs live in pinned host memory (the scene arrays in the reference's layout)
    pinned_scene = tuple(torch.from_numpy(np.ascontiguousarray(a, dtype=np.float32)).pin_memory().numpy()
                         for a in scene.as_tuple())
    last = {"ptr": None, "host": None}

    def device_step(camera=cam):
        if world == 1:
            ctx.render_bands_device(w, h, camera, BAND_ROWS, 0, 1, d_rgba8=d_frame.data_ptr(), stream=sptr)
        else:
            last["ptr"] = grp.render(camera, sptr)       # rank 0: the frame's slot, complete in stream order

    def e2e_step():
        # what a caller of the C ABI does per frame with HOST data: scene arrays in, frame out
        ctx.set_scene(*pinned_scene)
        if world == 1:
            ctx.render(w, h, cam, want=("rgba8",), out={"rgba8": h_frame})
            last["host"] = h_frame
        else:
            last["host"], _ = grp.render_host(cam)        # blocking; rank 0 returns when every rank has delivered

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def all_max(x):
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def timed_loop(step_fn, n):
        """n steps, each between two CUDA events on the stream, L2 flushed before each (not timed); device ms summed,
        max over ranks."""
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n)]
        barrier()
        wall0 = time.perf_counter()
        for i, (a, b) in enumerate(ev):
            flush.zero_()
            a.record(stream)
            step_fn(i)
            b.record(stream)
        barrier()
        wall = time.perf_counter() - wall0
        return all_max(sum(a.elapsed_time(b) for a, b in ev)) / n, wall * 1e3 / n

    def e2e_pass():
        """End to end through the C ABI with host buffers: per step the scene arrays go in (pinned host memory) and the
        frame comes out (pinned / shared page-locked host memory); wall clock over `steps` blocking calls, max over ranks."""
        ctx.set_option(capi.RTB_OPT_DEFER_SCENE_CHECK, 1)
        for _ in range(3):
            e2e_step()
        barrier()
        h2d0, d2h0 = ctx.transfer_bytes()
        t0 = time.perf_counter()
        for _ in range(steps):
            e2e_step()
        barrier()
        e2e_s = all_max(time.perf_counter() - t0)
        h2d1, d2h1 = ctx.transfer_bytes()
        ctx.set_option(capi.RTB_OPT_DEFER_SCENE_CHECK, 0)
        # bytes this rank moved; summed over the ranks (every rank uploads the scene, each downloads its bands)
        tb = torch.tensor([float(h2d1 - h2d0), float(d2h1 - d2h0)], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(tb)
        return dict(e2e_ms=e2e_s * 1e3 / steps, h2d=float(tb[0].item()) / steps, d2h=float(tb[1].item()) / steps)

    def timed_pass(mode):
        ctx.set_option(capi.RTB_OPT_TILE_CULL, mode)
        for _ in range(max(args.warmup, 3)):
            device_step()
        barrier()
        ctx.kernel_ms_avg(reset=True)
        sampler = ClockSampler(local_rank)
        sampler.start()
        n0 = ctx.launch_count()
        ms_per_step, wall_ms = timed_loop(lambda i: device_step(), steps)
        launches = ctx.launch_count() - n0
        k_avg_ms, k_n = ctx.kernel_ms_avg(reset=True)
        sampler.stop_flag = True
        sampler.join()
        out = dict(ms_per_step=ms_per_step, wall_ms_per_step=wall_ms, k_avg_ms=all_max(k_avg_ms), k_n=k_n, launches=launches,
                   clocks=sampler.result())
        # the same with the camera moving every step (orbit cameras): the FP64 record pack (and the dense kernel's row
        # table) are then part of every step; with the static camera above they are cached after the first frame
        if with_moving:
            cams = [hs.orbit_camera(w, h, 1 + (k % 119)) for k in range(min(steps, 12) + 2)]
            for c in cams[:2]:
                device_step(c)
            out["moving_ms_per_step"], _ = timed_loop(lambda i: device_step(cams[2 + i]), len(cams) - 2)
            device_step()                      # back to the reference camera
        barrier()
        out.update(e2e_pass())
        return out

    def frames_now():
        """(device frame, host frame) of the last steps as numpy on rank 0 (None elsewhere)."""
        torch.cuda.synchronize()
        if rank != 0:
            return None, None
        dev = d_frame if world == 1 else grp.frame_tensor(last["ptr"])
        return dev.cpu().numpy().copy(), np.array(last["host"], copy=True)

    dense = timed_pass(capi.TILE_CULL_DENSE)
    dev_dense, host_dense = frames_now()
    # End to end with the library's DEFAULT options (RTB_OPT_TILE_CULL = auto: what a caller of the C ABI gets without
    # asking; byte-identical frames): the headline e2e.  The dense kernel's e2e is kept beside it (e2e_dense).
    ctx.set_option(capi.RTB_OPT_TILE_CULL, capi.TILE_CULL_AUTO)
    default_e2e = e2e_pass()
    default_e2e["mode"] = ctx.get_option(capi.RTB_OPT_TILE_CULL)
    _, host_default = frames_now()
    if rank == 0:
        default_e2e["same_frame_as_dense"] = bool(np.array_equal(host_default, host_dense))
    # N > 1: the frame assembled from N GPUs' bands against the frame rank 0 renders on its own (dense kernel)
    match = None
    if world > 1:
        barrier()
        ok = {}
        if rank == 0:
            ctx.render_bands_device(w, h, cam, BAND_ROWS, 0, 1, d_rgba8=d_frame.data_ptr(), stream=sptr)
            torch.cuda.synchronize()
            single = d_frame.cpu().numpy()
            ok = {"peer_store_frame": bool(np.array_equal(dev_dense, single)), "shared_host_frame": bool(np.array_equal(host_dense, single))}
        barrier()
        match = ok
    binned = None
    if with_binned:
        # The same frame with the tile cull made once per frame (RTB_OPT_TILE_CULL = 2: bin_tiles_kernel +
        # raycast_binned_kernel, byte-identical output): reported beside the dense O(rays x primitives) numbers.
        binned = timed_pass(capi.TILE_CULL_BINNED)
        dev_b, host_b = frames_now()
        if rank == 0:
            binned["same_frame_as_dense"] = bool(np.array_equal(dev_b, dev_dense) and np.array_equal(host_b, host_dense))
            if match is not None:
                match["binned_frame"] = binned["same_frame_as_dense"] and match["peer_store_frame"]
    ctx.set_option(capi.RTB_OPT_TILE_CULL, capi.TILE_CULL_AUTO)
    barrier()
    if grp is not None:
        grp.close()
    torch.cuda.set_stream(torch.cuda.default_stream())
    del flush, d_frame
    return dict(key=key, w=w, h=h, ns=ns, nc=nc, seed=seed, scene=scene, my_rows=my_rows, dense=dense, binned=binned, match=match,
                host_frame=host_dense, default_e2e=default_e2e)


def run_ours(args):
    from opencl_ray_tracer_b200 import capi
    rank, world, local_rank = _setup_dist()
    if world != args.gpus and rank == 0:
        print("warning: --gpus %d but WORLD_SIZE %d; using WORLD_SIZE" % (args.gpus, world), file=sys.stderr)
    ctx = capi.Context(local_rank)                # raises if librtb.so / the device is missing
    main = measure_workload(args, args.workload, ctx, rank, world, local_rank, args.steps, not args.no_binned, True)
    c4 = None
    if args.workload == "c3" and not args.no_c4:
        # BASELINE config 4 (7680x4320, 4096 primitives: the config named for sharding) in every line, fewer steps
        c4 = measure_workload(args, "c4", ctx, rank, world, local_rank, max(3, min(args.steps, 10)), not args.no_binned, False)
    return dict(rank=rank, world=world, ctx=ctx, main=main, c4=c4)


def run_sequence(args):
    """BASELINE config 5: the 120-frame orbiting-camera sequence at C3's size.  One step = the whole
    sequence, frames device-resident (N = 1: rtb_render_sequence_device, consecutive frames on two alternating
    streams of the context so that a frame's ragged end overlaps the next frame's start; N > 1: every frame sharded as
    row bands, every rank's kernels storing into rank 0's sequence buffer over NVLink, ONE fence at the
    end).  e2e = every frame delivered into a pinned / shared host frame (rtb_render, rtb_render_bands_host)."""
    import torch
    import torch.distributed as dist
    from opencl_ray_tracer_b200 import bands, capi, host_scene as hs
    rank, world, local_rank = _setup_dist()
    base, n_frames, deg = SEQUENCE[args.workload]
    w, h, ns, nc, seed = WORKLOADS[base]
    scene = _broadcast_scene(hs.random_scene(w, h, ns, nc, seed), rank, world)
    ctx = capi.Context(local_rank)
    ctx.set_option(capi.RTB_OPT_TILE_CULL, capi.TILE_CULL_DENSE)
    ctx.set_scene(*scene.as_tuple())
    cams = [hs.orbit_camera(w, h, k) for k in range(n_frames)]
    flat = np.stack([np.concatenate(c) for c in cams])
    stream = torch.cuda.Stream()
    torch.cuda.set_stream(stream)
    sptr = stream.cuda_stream
    # N > 1: a ring of 8 frame slots on rank 0 (the ranks run up to 7 frames ahead of rank 0's stream)
    grp = capi.Group(ctx, _group_name("c5", rank, world), rank, world, w, h, BAND_ROWS, 8) if world > 1 else None
    # N > 1, --lanes 2 (default): a SECOND context + group + stream per rank; frames alternate between the two, so a
    # frame's ragged end, its NVLink drain into rank 0 and its stream operations overlap the next frame's kernel (what
    # rtb_render_sequence_device does inside one context for N = 1).  Each lane is the plain public API.
    lanes = [(ctx, grp, stream)]
    if world > 1 and args.lanes >= 2:
        ctx2 = capi.Context(local_rank)
        ctx2.set_option(capi.RTB_OPT_TILE_CULL, capi.TILE_CULL_DENSE)
        ctx2.set_scene(*scene.as_tuple())
        lanes.append((ctx2, capi.Group(ctx2, _group_name("c5b", rank, world), rank, world, w, h, BAND_ROWS, 8), torch.cuda.Stream()))
    d_out = torch.empty((n_frames, h, w, 4), dtype=torch.uint8, device="cuda") if world == 1 else None
    h_frame = torch.empty((h, w, 4), dtype=torch.uint8).pin_memory().numpy() if world == 1 else None
    token = torch.zeros(1, dtype=torch.int32, device="cuda")

    def device_step():
        if world == 1:
            ctx.render_sequence_device(w, h, flat, d_out.data_ptr())     # blocking
            return
        nl = len(lanes)
        for k, (_, g, st) in enumerate(lanes):   # the frames' records (this rank's bands) are packed ahead on each context's side stream
            g.prepack(flat[k::nl], st.cuda_stream)
        for f in range(n_frames):
            _, g, st = lanes[f % nl]
            g.render(cams[f], st.cuda_stream)    # one C call per frame and rank; the lane's stream holds the frame's completion
        for _, _, st in lanes:
            st.synchronize()

    def e2e_step():
        for f in range(n_frames):
            if world == 1:
                ctx.render(w, h, cams[f], want=("rgba8",), out={"rgba8": h_frame})
            else:
                grp.render_host(cams[f])

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            fn()
        barrier()
        t = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    steps = max(1, args.steps)
    for _ in range(max(1, min(args.warmup, 3))):
        device_step()
    for c, _, _ in lanes:
        c.kernel_ms_avg(reset=True)
    sampler = ClockSampler(local_rank)
    sampler.start()
    n_launch0 = sum(c.launch_count() for c, _, _ in lanes)
    dev_s = timed(device_step, steps)
    launches = sum(c.launch_count() for c, _, _ in lanes) - n_launch0
    sampler.stop_flag = True
    sampler.join()
    k_pairs = [c.kernel_ms_avg(reset=True) for c, _, _ in lanes]
    k_n = sum(n for _, n in k_pairs)
    k_avg_ms = sum(a * n for a, n in k_pairs) / k_n if k_n else 0.0
    # the same sequence with the binned tile cull (RTB_OPT_TILE_CULL = 2, byte-identical frames), device-resident
    binned_s = None
    if not args.no_binned:
        for c, _, _ in lanes:
            c.set_option(capi.RTB_OPT_TILE_CULL, 2)
        device_step()
        binned_s = timed(device_step, steps) / steps
        for c, _, _ in lanes:
            c.set_option(capi.RTB_OPT_TILE_CULL, capi.TILE_CULL_DENSE)
    ctx.set_option(capi.RTB_OPT_TILE_CULL, capi.TILE_CULL_AUTO)   # e2e: the library's default options, as in the single-frame lines
    e2e_step()
    e2e_steps = max(1, steps // 4)
    h2d0, d2h0 = ctx.transfer_bytes()
    e2e_s = timed(e2e_step, e2e_steps)
    h2d1, d2h1 = ctx.transfer_bytes()
    tb = torch.tensor([float(h2d1 - h2d0), float(d2h1 - d2h0)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tb)
    if rank == 0:
        n_px = w * h
        s_step = dev_s / steps
        sms = ctx.device_info()["sm_count"]
        my_rows = ctx.band_local_rows(h, BAND_ROWS, 0, world)
        px, tile_w = capi.debug_grid_shape(w, my_rows, sms)[:2]
        flop_per_test = FLOP_PER_TEST
        probe = fp32_probe_tflops()
        peak_nominal = sms * 128 * 2 * (sampler.result()["sm_max_mhz"] or 1965.0) * 1e6 / 1e12
        achieved = my_rows * w * (ns + nc) * flop_per_test / (k_avg_ms * 1e-3) / 1e12 if k_avg_ms > 0 else 0.0
        peak = probe if probe else peak_nominal
        line = {
            "metric": "Mrays/s", "value": n_frames * n_px / s_step / 1e6, "unit": "Mrays/s", "n_gpus": world,
            "steps": steps, "warmup": max(1, min(args.warmup, 3)), "ms_per_step": s_step * 1e3, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": dict(bench_config(args.workload, world),
                           l2="each step writes %d frames = %.1f GB of output (larger than L2); a step is timed as a whole (wall clock around blocking calls, max over ranks)" % (
                               n_frames, n_frames * 4 * n_px / 1e9),
                           records="moving camera: every frame's records and row table are packed (ahead, on a side stream)",
                           lanes=("%d contexts (+ groups, streams) per rank, frames alternate between them" % len(lanes)) if world > 1 else
                                 "one context; rtb_render_sequence_device alternates two streams inside it"),
            "frames_per_s": n_frames / s_step, "ms_per_frame": s_step * 1e3 / n_frames,
            "tests_per_s": n_frames * n_px * (ns + nc) / s_step,
            "clocks": sampler.result(),
            "e2e": {"value": n_frames * n_px * e2e_steps / e2e_s / 1e6, "unit": "Mrays/s",
                    "frames_per_s": n_frames * e2e_steps / e2e_s,
                    "h2d_bytes_per_step": float(tb[0].item()) / e2e_steps, "d2h_bytes_per_step": float(tb[1].item()) / e2e_steps,
                    "ms_per_step": e2e_s * 1e3 / e2e_steps,
                    "what": "every frame rendered with the library's default options (RTB_OPT_TILE_CULL = auto) and delivered into a pinned (N = 1) / shared page-locked (N > 1) host frame by one blocking call per frame; the scene stays resident, the step's input is the camera sequence"},
            "gpu_launches": launches,
            "roofline": {"bound": "fp32", "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak if peak else None,
                         "traffic": None, "kernel": "raycast_dense_kernel", "kernel_ms": k_avg_ms, "kernel_launches_timed": k_n,
                         "flop_per_test": flop_per_test, "rays_per_thread": px,
                         "peak_source": ("fp32_issue_probe (packed FFMA2 chains) run on this GPU just now" if probe
                                         else "nominal SMs x 128 lanes x 2 x max SM clock")},
        }
        if binned_s:
            line["binned"] = {"what": "the same sequence rendered with RTB_OPT_TILE_CULL = 2 (bin_tiles_kernel + raycast_binned_kernel per frame, "
                                      "byte-identical frames), device-resident", "frames_per_s": n_frames / binned_s,
                              "ms_per_frame": binned_s * 1e3 / n_frames, "value": n_frames * n_px / binned_s / 1e6, "unit": "Mrays/s"}
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        for _, g, _ in lanes:
            g.close()
        dist.destroy_process_group()


# ---- reference arm -------------------------------------------------------------------------------
_REF = {}


def _ref_worker_init(sph_o, sph_r, sph_c, cube_c, recipes, kind):
    from oracle.oracle import PortOracle, RefOracle, Scene
    if kind == "reference" and recipes is None:
        ref = RefOracle()
        ref.build_scene(1)             # c1: the reference's own createScene1 (MainState.cpp:419-462)
        _REF["ref"] = ref
    elif kind == "reference":
        ref = RefOracle()
        ref.clear_scene()
        for o, r, c in zip(sph_o, sph_r, sph_c):
            ref.add_sphere(o, r, c)
        for c, q in zip(cube_c, recipes):  # same calls as createScene3's loop body, MainState.cpp:618-638
            ref.add_cube_ops(c, [("scale", (q[0],) * 3), ("rotate", (0, 0, q[1])), ("rotate", (0, q[2], 0)),
                                 ("rotate", (q[3], 0, 0)), ("translate", (q[4], q[5], q[6]))])
        _REF["ref"] = ref
    else:
        _REF["port"] = PortOracle()


def _ref_worker_rows(job):
    w, h, segs, scene_tuple = job
    d = np.array([0, 0, -1, -1], np.float32)
    n = 0
    for (y, x0, x1) in segs:
        o = np.zeros((x1 - x0, 4), np.float32)
        o[:, 0], o[:, 1], o[:, 3] = np.arange(x0, x1, dtype=np.float32), float(y), 1.0
        if "ref" in _REF:
            _REF["ref"].render(o, d)   # the reference's own executeRayTracerCPU loop body + collide
        else:
            from oracle.oracle import Scene
            sc = Scene(*scene_tuple)
            _REF["port"].render(sc, x1 - x0, 1, ray_origins=o, ray_dir=d, faithful=True, threads=1)
        n += x1 - x0
    return n


def _time_opencl_kernel_on_host(scene, w, h, target_seconds=8.0):
    """The reference's OTHER path, timed beside its CPU path: the kernel source rayTracer.cl compiled
    as C++ (oracle/_ref/libraytracer_cl_ref.so; there is no OpenCL ICD on this box), one call per work
    item, OpenMP over the work items on all host cores, on a bounded sample of rows of the same frame."""
    from oracle.oracle import ClRefOracle, Scene, cl_ref_available
    if not cl_ref_available():
        return None
    cl = ClRefOracle()
    sc = Scene(*scene.as_tuple())
    d = np.array([0, 0, -1, -1], np.float32)

    def rows(ys):
        o = np.zeros((len(ys), w, 4), np.float32)
        o[..., 0] = np.arange(w, dtype=np.float32)[None, :]
        o[..., 1] = np.asarray(ys, np.float32)[:, None]
        o[..., 3] = 1.0
        t0 = time.perf_counter()
        cl.render(sc, o.reshape(-1, 4), d)
        return time.perf_counter() - t0
    probe = list(range(0, h, max(1, h // 8)))[:8]
    dt = rows(probe)
    n = int(max(len(probe), min(h, len(probe) * target_seconds / max(dt, 1e-6))))
    ys = list(range(0, h, max(1, h // n)))[:n]
    dt = rows(ys)
    return {"value": len(ys) * w / dt / 1e6, "unit": "Mrays/s", "cores": os.cpu_count() or 1,
            "kind": "reference: rayTracer.cl compiled as C++ for the host (OpenCL ICD unavailable on this box)",
            "sample": "%d of %d rows of the same %dx%d frame, %.1f s" % (len(ys), h, w, h, dt)}


class _PlainScene:
    """Scene arrays of the reference arm (no library of this repo behind them)."""

    def __init__(self, z):
        self.sph_o, self.sph_r, self.sph_c, self.cube_v, self.cube_c = (z[k] for k in ("sph_o", "sph_r", "sph_c", "cube_v", "cube_c"))
        self.cube_recipes = z["recipes"] if len(z["recipes"]) == len(self.cube_c) and len(self.cube_c) else None
        self.ns, self.nc = len(self.sph_r), len(self.cube_c)

    def as_tuple(self):
        return self.sph_o, self.sph_r, self.sph_c, self.cube_v, self.cube_c


def _scene_from_child_process(key):
    """The workload's scene arrays for the reference arm.  They come from the repo's host mirror (librtb_host.so), which
    must not be loaded into the process that times the reference: a child python builds them and writes an .npz."""
    import subprocess
    import tempfile
    path = os.path.join(tempfile.mkdtemp(prefix="rtb_ref_scene_"), key + ".npz")
    code = ("import sys; sys.path.insert(0, %r); import numpy as np, bench; s = bench.make_scene(%r); "
            "np.savez(%r, sph_o=s.sph_o, sph_r=s.sph_r, sph_c=s.sph_c, cube_v=s.cube_v, cube_c=s.cube_c, "
            "recipes=(s.cube_recipes if s.cube_recipes is not None else np.zeros((0, 7), np.float32)))" % (ROOT, key, path))
    subprocess.run([sys.executable, "-c", code], check=True)
    with np.load(path) as z:
        sc = _PlainScene({k: z[k] for k in z.files})
    os.remove(path)
    return sc


def _loaded_libraries(pattern):
    """Shared objects mapped into this process whose path contains `pattern` (from /proc/self/maps)."""
    try:
        return sorted({ln.split()[-1] for ln in open("/proc/self/maps") if pattern in ln and ".so" in ln})
    except OSError:
        return []


def _ref_worker_libs(_):
    return _loaded_libraries(ROOT)


def run_reference(args):
    """The reference's CPU implementation of the path on this box's host cores: the per-ray code is
    the reference's own (oracle/_ref: MainState::collide / intersectTri / intersectSphere compiled
    from its sources) when that library was built, else the faithful mode of the oracle port.  The
    reference is single-threaded; to use every host core the rays of the sample are spread over one
    worker process per core, each running the unmodified per-ray loop.  Each step is a bounded
    sample (row segments spread over the frame) sized so that the whole run takes about 90 s."""
    import multiprocessing as mp
    from oracle.oracle import ref_available
    key = SEQUENCE[args.workload][0] if args.workload in SEQUENCE else args.workload
    w, h = WORKLOADS[key][:2]
    scene = _scene_from_child_process(key)
    ns, nc = scene.ns, scene.nc
    kind = "reference" if (ref_available() and (scene.cube_recipes is not None or key == "c1")) else "port"
    cores = os.cpu_count() or 1
    init = (scene.sph_o, scene.sph_r, scene.sph_c, scene.cube_c, scene.cube_recipes, kind)
    ctxm = mp.get_context("fork")
    with ctxm.Pool(cores, initializer=_ref_worker_init, initargs=init) as pool:
        worker_libs = sorted({p for libs in pool.map(_ref_worker_libs, range(cores)) for p in libs})
        def step(k, seg_len):
            ys = [int((i + 0.5) * h / cores + 7 * k) % h for i in range(cores)]   # one row per worker, spread
            x0 = (k * seg_len) % max(1, w - seg_len + 1)
            jobs = [(w, h, [(ys[i], x0, x0 + seg_len)], scene.as_tuple() if kind == "port" else None)
                    for i in range(cores)]
            t0 = time.perf_counter()
            rays = sum(pool.map(_ref_worker_rows, jobs))
            return rays, time.perf_counter() - t0
        probe_len = min(w, 256)
        _, s = step(0, probe_len)                       # untimed probe: rays/s per worker
        rate = probe_len / max(s, 1e-6)
        budget = 90.0 / max(1, args.steps + args.warmup)
        seg_len = int(max(64, min(w, rate * budget)))
        for k in range(args.warmup):
            step(k, seg_len)
        tot_rays, tot_s = 0, 0.0
        for k in range(args.steps):
            r, s = step(args.warmup + k, seg_len)
            tot_rays += r
            tot_s += s
    opencl_path = _time_opencl_kernel_on_host(scene, w, h)
    mrays = tot_rays / tot_s / 1e6
    sample = ("%d row segments of %d rays per step (one per worker process, rows spread over the frame: a bounded sample, not a "
              "whole frame), %d steps, of the same %dx%d frame; the reference is single-threaded -- one worker process per host "
              "core each runs its unmodified per-ray loop, which flatters it by the core count") % (cores, seg_len, args.steps, w, h)
    line = {
        "impl": "reference", "metric": "Mrays/s", "value": mrays, "unit": "Mrays/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": tot_s * 1e3 / args.steps,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32 (f64 triangle test)",
        "data": "synthetic",
        "config": bench_config(args.workload, args.gpus),     # the same dict as our arm's line (the frame and scene are the same;
                                                              # its parallelism / l2 / records entries describe our arm)
        "reference_sample_of": (None if args.workload in WORKLOADS else "frame 0 of the orbit"),
        "tests_per_s": mrays * 1e6 * (ns + nc),
        "libraries_loaded_by_the_workers": worker_libs,
        "cpu_baseline": {"value": mrays, "unit": "Mrays/s", "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": mrays, "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    if opencl_path:
        line["opencl_path_on_host"] = opencl_path
    print(json.dumps(line))


def pair_rate_probe():
    """profiles/microbench/pair_rate_probe: scheduler clocks per (FFMA2 + FMNMX3) pair on this GPU, alone and with the
    loop's other instructions mixed in -- the per-ray-pair, per-primitive cost that bounds the dense kernel."""
    import subprocess
    exe = os.path.join(ROOT, "profiles", "microbench", "pair_rate_probe")
    if not os.path.isfile(exe):
        return None
    try:
        out = subprocess.run([exe], capture_output=True, text=True, timeout=60).stdout
        res = {}
        for line in out.splitlines():
            j = json.loads(line)
            if "clocks_per_ffma2_fmnmx3_pair" in j:
                res["with_loop_overhead" if "LDS" in j["mix"] else "pair_only"] = j["clocks_per_ffma2_fmnmx3_pair"]
            elif j.get("mix") == "8xFFMA2":
                res["ffma2"] = j["clocks_per_instr"]
            elif j.get("mix", "").startswith("8xFMNMX3"):
                res["fmnmx3"] = j["clocks_per_instr"]
        return res or None
    except Exception:
        return None


def workload_record(args, r, world, ctx, peaks, peaks_src, probe, pair):
    """The JSON fields of one measured workload (measure_workload's result `r`)."""
    from opencl_ray_tracer_b200 import capi
    w, h, ns, nc = r["w"], r["h"], r["ns"], r["nc"]
    n_px = w * h
    d = r["dense"]
    steps = None
    info = ctx.device_info()
    sms = info["sm_count"]
    sm_max = (d["clocks"]["sm_max_mhz"] or peaks.get("sm_max_mhz", 1965.0))
    peak_nominal = sms * 128 * 2 * sm_max * 1e6 / 1e12          # TFLOP/s at the max SM clock
    px, tile_w = capi.debug_grid_shape(w, r["my_rows"], sms)[:2]
    k_s = d["k_avg_ms"] * 1e-3
    tests_launch = r["my_rows"] * w * (nc + ns)
    achieved = tests_launch * FLOP_PER_TEST / k_s / 1e12 if k_s > 0 else 0.0
    peak = probe if probe else peak_nominal
    pair_clk = (pair or {}).get("pair_only") or PAIR_CLOCKS_FALLBACK
    disp_peak = sms * 4 * 32 * sm_max * 1e6 * 2.0 / pair_clk      # tests/s: 2 rays per pair, 32 lanes, 4 schedulers per SM
    rec = {
        "value": n_px / (d["ms_per_step"] * 1e-3) / 1e6, "unit": "Mrays/s", "ms_per_step": d["ms_per_step"],
        "kernel_ms": d["k_avg_ms"], "step_minus_kernel_ms": d["ms_per_step"] - d["k_avg_ms"],
        "tests_per_s": n_px * (ns + nc) / (d["ms_per_step"] * 1e-3),
        "wall_ms_per_step_incl_flush": d["wall_ms_per_step"], "gpu_launches": d["launches"],
        "moving_camera_ms_per_step": d.get("moving_ms_per_step"),
        "e2e": {"value": n_px / (r["default_e2e"]["e2e_ms"] * 1e-3) / 1e6, "unit": "Mrays/s", "ms_per_step": r["default_e2e"]["e2e_ms"],
                "h2d_bytes_per_step": r["default_e2e"]["h2d"], "d2h_bytes_per_step": r["default_e2e"]["d2h"],
                "tile_cull_mode": r["default_e2e"]["mode"], "same_frame_as_dense": r["default_e2e"].get("same_frame_as_dense")},
        "e2e_dense": {"value": n_px / (d["e2e_ms"] * 1e-3) / 1e6, "unit": "Mrays/s", "ms_per_step": d["e2e_ms"],
                      "h2d_bytes_per_step": d["h2d"], "d2h_bytes_per_step": d["d2h"]},
        "roofline": {"bound": "fp32", "achieved": achieved, "peak": peak, "unit": "TFLOP/s",
                     "frac": achieved / peak if peak else None,
                     "traffic": ncu_traffic_bytes(r["key"]) if world == 1 else None,
                     "traffic_unit": "bytes of DRAM traffic per launch (ncu capture in profiles/, not this run)",
                     "kernel": "raycast_dense_kernel", "kernel_ms": d["k_avg_ms"], "kernel_launches_timed": d["k_n"],
                     "flop_per_test": FLOP_PER_TEST, "rays_per_thread": px, "tile_width": tile_w,
                     "peak_source": ("fp32_issue_probe (packed FFMA2 chains) run on this GPU just now" if probe
                                     else "nominal SMs x 128 lanes x 2 x max SM clock"),
                     "peak_nominal_tflops": peak_nominal, "frac_of_nominal": achieved / peak_nominal,
                     "north_star_target_frac": 0.60,
                     "note": "the >= 60 % FP32-FMA target of BASELINE.md is NOT met: the shipped formulation needs ONE FMA per "
                             "ray-primitive test and its other instruction (the 3-input FMNMX of the running minimum) does not "
                             "run in the shadow of the FMAs -- see `dispatch`",
                     "algorithmic_flops_at_baseline_constants": {
                         "what": "the same tests/s priced at BASELINE.md's per-ray world-space formulation (27 flop per cube test, "
                                 "18 per sphere test) -- arithmetic this kernel does not execute; for comparison only",
                         "tflops": (r["my_rows"] * w * (27.0 * nc + 18.0 * ns) / k_s / 1e12) if k_s > 0 else None},
                     "dispatch": {"what": "tests per second against the measured issue cost of the loop body: one FFMA2 and one "
                                          "FMNMX3 per ray pair and primitive = %.2f scheduler clocks (pair_rate_probe on this GPU%s); "
                                          "peak = SMs x 4 schedulers x 32 lanes x 2 rays x max SM clock / that" % (
                                              pair_clk, "" if pair else ": probe missing, round-2 measurement used"),
                                  "clocks_per_pair": pair_clk, "probe": pair,
                                  "achieved_tests_per_s": tests_launch / k_s if k_s > 0 else None,
                                  "peak_tests_per_s": disp_peak,
                                  "frac": (tests_launch / k_s) / disp_peak if k_s > 0 else None},
                     "framebuffer_write_gbs": r["my_rows"] * w * 4 / k_s / 1e9 if k_s > 0 else None,
                     "hbm_peak_gbs": peaks.get("hbm_gbs"), "hbm_peak_source": peaks_src},
        "clocks": d["clocks"],
    }
    if r["binned"]:
        bn = r["binned"]
        bk_s = bn["k_avg_ms"] * 1e-3
        frame_bytes = r["my_rows"] * w * 4
        rec["binned"] = {
            "what": "the same frame, byte for byte, with RTB_OPT_TILE_CULL = 2 (what the default, auto, picks from 64 primitives on): "
                    "bin_tiles_kernel decides once per frame which primitives a 64x32-pixel tile can see, raycast_binned_kernel runs "
                    "the per-ray tests for those only (not O(rays x primitives): no tests/s or FP32 fraction is claimed for it)",
            "value": n_px / (bn["ms_per_step"] * 1e-3) / 1e6, "unit": "Mrays/s", "ms_per_step": bn["ms_per_step"],
            "kernels_ms": bn["k_avg_ms"], "step_minus_kernel_ms": bn["ms_per_step"] - bn["k_avg_ms"], "gpu_launches": bn["launches"],
            "moving_camera_ms_per_step": bn.get("moving_ms_per_step"), "same_frame_as_dense": bn.get("same_frame_as_dense"),
            "e2e": {"value": n_px / (bn["e2e_ms"] * 1e-3) / 1e6, "unit": "Mrays/s", "ms_per_step": bn["e2e_ms"],
                    "h2d_bytes_per_step": bn["h2d"], "d2h_bytes_per_step": bn["d2h"]},
            "roofline": {"bound": "hbm", "achieved": frame_bytes / bk_s / 1e9 if bk_s > 0 else None, "peak": peaks.get("hbm_gbs"),
                         "unit": "GB/s", "frac": (frame_bytes / bk_s / 1e9 / peaks["hbm_gbs"]) if (bk_s > 0 and peaks.get("hbm_gbs")) else None,
                         "traffic": ncu_traffic_bytes(r["key"], "binned") if world == 1 else None,
                         "traffic_unit": "bytes of DRAM traffic per frame (ncu capture in profiles/, not this run): the frame stays in the 126 MB L2",
                         "kernel": "bin_tiles_kernel + raycast_binned_kernel",
                         "algorithmic_bytes": "the frame write, 4 B per ray (this rank's rows)", "peak_source": peaks_src},
            "clocks": bn["clocks"],
        }
    if r["match"] is not None:
        rec["frame_matches_single_gpu"] = r["match"]
    return rec


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    if args.impl == "reference":
        if rank == 0:
            run_reference(args)
        return
    if args.workload in SEQUENCE:
        run_sequence(args)
        return
    r = run_ours(args)
    import torch.distributed as dist
    world = r["world"]
    if r["rank"] == 0:
        peaks, peaks_src = load_peaks()
        probe = fp32_probe_tflops()
        pair = pair_rate_probe()
        m = r["main"]
        rec = workload_record(args, m, world, r["ctx"], peaks, peaks_src, probe, pair)
        line = {
            "metric": "Mrays/s", "value": rec["value"], "unit": "Mrays/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": rec["ms_per_step"], "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": bench_config(args.workload, world),
        }
        for k in ("tests_per_s", "kernel_ms", "step_minus_kernel_ms", "wall_ms_per_step_incl_flush", "moving_camera_ms_per_step", "clocks"):
            line[k] = rec[k]
        how = ("rtb_set_scene (arrays staged in pinned memory, H2D, boxes recovered on the device; RTB_OPT_DEFER_SCENE_CHECK = 1: no wait "
               "for the device) + rtb_render (device record pack, kernels, frame copied to pinned host memory)" if world == 1 else
               "per rank and frame: rtb_set_scene + rtb_group_render_host -- each GPU copies its own bands over its own PCIe link into "
               "ONE frame in shared, page-locked host memory; completion and release through counters in the ranks' shared control "
               "block (no collective)")
        line["e2e"] = dict(rec["e2e"], what="the call a user of the C ABI makes, with the library's DEFAULT options (RTB_OPT_TILE_CULL = auto: the "
                                            "binned kernels from 64 primitives on; the frame is byte-identical to the dense kernel's): " + how)
        line["e2e_dense"] = dict(rec["e2e_dense"], what="the same with RTB_OPT_TILE_CULL = 0 (the dense O(rays x primitives) kernel that `value` and "
                                                        "`roofline` are quoted on; one launch, finished row chunks copied out while it runs)")
        line["gpu_launches"] = rec["gpu_launches"]
        line["roofline"] = rec["roofline"]
        if "binned" in rec:
            line["binned"] = rec["binned"]
        if "frame_matches_single_gpu" in rec:
            line["frame_matches_single_gpu"] = rec["frame_matches_single_gpu"]
        if r["c4"] is not None:
            c4 = workload_record(args, r["c4"], world, r["ctx"], peaks, peaks_src, probe, pair)
            sub = {"what": "BASELINE config 4 in the same run: " + workload_name("c4"), "steps": max(3, min(args.steps, 10))}
            for k in ("value", "unit", "ms_per_step", "kernel_ms", "step_minus_kernel_ms", "tests_per_s", "e2e", "e2e_dense", "gpu_launches"):
                sub[k] = c4[k]
            sub["roofline"] = {k: c4["roofline"][k] for k in ("bound", "achieved", "peak", "unit", "frac", "frac_of_nominal", "kernel_ms", "dispatch")}
            if "binned" in c4:
                sub["binned"] = {k: c4["binned"][k] for k in ("value", "unit", "ms_per_step", "kernels_ms", "same_frame_as_dense", "e2e")}
            if "frame_matches_single_gpu" in c4:
                sub["frame_matches_single_gpu"] = c4["frame_matches_single_gpu"]
            line["c4"] = sub
        if world == 1 and not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline_port(m["scene"], m["w"], m["h"], args.cpu_seconds)
            if args.workload in ("c1", "c2"):   # full-frame single-thread timing of the reference's own loop
                line["cpu_faithful"] = cpu_faithful_frame(args.workload, m["scene"], m["w"], m["h"], m["host_frame"])
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
