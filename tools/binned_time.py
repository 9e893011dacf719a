"""Kernel time of a binned frame (bin_tiles_kernel + raycast_binned_kernel, L2 flushed before each) for C3 / C4 / C2 on one GPU, with the frame's
hash; RTB_LIB selects another build for an A/B (e.g. a copy of the previous librtb.so)."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from opencl_ray_tracer_b200 import capi, host_scene as hs
import hashlib
ctx = capi.Context(0)
ctx.set_option(capi.RTB_OPT_TILE_CULL, 2)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
stream = torch.cuda.Stream()
for (w, h, ns, nc, seed) in ((3840, 2160, 512, 512, 3), (7680, 4320, 2048, 2048, 4), (1920, 1080, 64, 64, 2)):
    ctx.set_scene(*hs.random_scene(w, h, ns, nc, seed).as_tuple())
    buf = torch.zeros((h, w, 4), dtype=torch.uint8, device="cuda")
    with torch.cuda.stream(stream):
        for _ in range(5):
            ctx.render_bands_device(w, h, capi.REFERENCE_CAMERA, 16, 0, 1, d_rgba8=buf.data_ptr(), stream=stream.cuda_stream)
        stream.synchronize(); ctx.kernel_ms_avg(reset=True)
        for _ in range(30):
            flush.zero_()
            ctx.render_bands_device(w, h, capi.REFERENCE_CAMERA, 16, 0, 1, d_rgba8=buf.data_ptr(), stream=stream.cuda_stream)
        stream.synchronize()
    ms, n = ctx.kernel_ms_avg(reset=True)
    print(json.dumps({"lib": os.path.basename(os.environ.get("RTB_LIB", "librtb.so")), "frame": "%dx%d/%d" % (w, h, ns + nc), "binned_kernels_ms": round(ms, 5), "sha1": hashlib.sha1(buf.cpu().numpy().tobytes()).hexdigest()[:12]}), flush=True)
    del buf
ctx.close()
