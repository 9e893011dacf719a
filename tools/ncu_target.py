"""Small fixed launch sequence for ncu (tools/gpu_ncu.sh): C3's scene, L2 flushed before every frame.
  2 x dense kernel, whole frame (2160 rows)          raycast_dense_kernel
  2 x dense kernel, one rank's bands at world 8      raycast_dense_kernel (272 rows)
  2 x dense kernel, one rank's bands at world 2      raycast_dense_kernel (1088 rows)
  2 x binned frame                                   bin_tiles_kernel + raycast_binned_kernel
With "c4": the same for BASELINE config 4 (whole frame and world 8 only)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from opencl_ray_tracer_b200 import capi, host_scene as hs  # noqa: E402

c4 = "c4" in sys.argv[1:]
w, h, ns, nc, seed = (7680, 4320, 2048, 2048, 4) if c4 else (3840, 2160, 512, 512, 3)
ctx = capi.Context(0)
ctx.set_scene(*hs.random_scene(w, h, ns, nc, seed).as_tuple())
buf = torch.zeros((h, w, 4), dtype=torch.uint8, device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
stream = torch.cuda.Stream()
with torch.cuda.stream(stream):
    for mode, world in ((0, 1), (0, 8)) + (() if c4 else ((0, 2), (2, 1))):
        ctx.set_option(capi.RTB_OPT_TILE_CULL, mode)
        for _ in range(2):
            flush.zero_()
            ctx.render_bands_device(w, h, capi.REFERENCE_CAMERA, 16, 0, world, d_rgba8=buf.data_ptr(), stream=stream.cuda_stream)
        stream.synchronize()
ms, n = ctx.kernel_ms_avg(reset=True)
print("launch groups timed: %d, avg %.4f ms" % (n, ms))
ctx.close()
