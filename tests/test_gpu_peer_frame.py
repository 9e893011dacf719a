"""Multi-GPU frame path without a gather: every rank's grid kernel stores its interleaved row bands
straight into rank 0's frame through a CUDA-IPC mapping (NVLink / NVSwitch peer memory on a
multi-GPU box).  The round-end GPU tests run on ONE B200, so the two ranks here are two processes
on cuda:0 (IPC between processes works on one device; NCCL does not, hence gloo for the 64-byte
handle and the completion token).  Checked: the shared frame equals the whole-frame render."""
import os
import socket

import numpy as np
import pytest

from parity import ROOT  # noqa: F401  (puts the repo root on sys.path, also for the spawned ranks)

pytestmark = pytest.mark.gpu


def _worker(rank, world, port, w, h, band_rows, q):
    import torch
    import torch.distributed as dist
    from opencl_ray_tracer_b200 import bands, capi, host_scene as hs
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        ctx = capi.Context(0)
        s = hs.random_scene(w, h, 40, 40, 31)
        ctx.set_scene(*s.as_tuple())
        pf = bands.PeerFrame(dist, ctx, w, h, rank, world)
        if not pf.ok:
            if rank == 0:
                q.put("ipc-unavailable")
            return
        pf.render(capi.REFERENCE_CAMERA, band_rows)   # stream 0 -> synchronous in the context
        torch.cuda.synchronize()
        pf.fence()                                    # gloo: every rank's kernel has completed
        ok = True
        if rank == 0:
            whole = ctx.render(w, h, want=("rgba8",))["rgba8"]
            got = pf.as_tensor().cpu().numpy()
            ok = bool(np.array_equal(got, whole))
            pf.as_tensor().zero_()
            torch.cuda.synchronize()
        dist.barrier()
        # the same frame with the signal fence: each launch's last CTA bumps rank 0's signal word and
        # rank 0 waits on it with a stream memory operation -- no collective before the frame is read
        for _ in range(2):
            pf.render(capi.REFERENCE_CAMERA, band_rows, signal=True)
        if rank == 0:
            pf.wait_all()                             # blocks until 2 frames x world signals have arrived
            got = pf.as_tensor().cpu().numpy()
            ok = ok and bool(np.array_equal(got, whole)) and pf.signals_expected == 2 * world
            q.put(ok)
        dist.barrier()
        pf.close()
        ctx.close()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,w,h,band_rows", [(2, 640, 480, 16), (3, 333, 203, 4)])
def test_bands_written_into_rank0_frame_over_ipc(built_libs, world, w, h, band_rows):
    import torch.multiprocessing as mp
    sk = socket.socket()
    sk.bind(("127.0.0.1", 0))
    port = sk.getsockname()[1]
    sk.close()
    mctx = mp.get_context("spawn")
    q = mctx.Queue()
    procs = [mctx.Process(target=_worker, args=(r, world, port, w, h, band_rows, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = q.get(timeout=300)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    if res == "ipc-unavailable":
        pytest.skip("CUDA IPC is not available in this container")
    assert res is True


def _host_worker(rank, world, port, w, h, band_rows, q):
    import torch.distributed as dist
    from opencl_ray_tracer_b200 import bands, capi, host_scene as hs
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        ctx = capi.Context(0)
        s = hs.random_scene(w, h, 40, 40, 31)
        ctx.set_scene(*s.as_tuple())
        shf = bands.SharedHostFrame(dist, w, h, rank, world)
        if rank == 0:
            shf.frame[:] = 7
        dist.barrier()
        for _ in range(2):                      # twice: the progress flags carry a new epoch per frame
            kms = shf.render(ctx, capi.REFERENCE_CAMERA, band_rows)
            assert kms > 0
        dist.barrier()                          # every rank's bands are in the shared frame
        if rank == 0:
            whole = ctx.render(w, h, want=("rgba8",))["rgba8"]
            q.put(bool(np.array_equal(shf.frame, whole)))
        dist.barrier()
        shf.close()
        ctx.close()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,w,h,band_rows", [(2, 1920, 1080, 16), (3, 333, 203, 4), (1, 640, 480, 16)])
def test_bands_delivered_into_shared_host_frame(built_libs, world, w, h, band_rows):
    """rtb_render_bands_host: every rank copies its bands (finished chunks first, while its launch is
    still running) into ONE frame in shared host memory; the frame equals the whole-frame render."""
    import torch.multiprocessing as mp
    sk = socket.socket()
    sk.bind(("127.0.0.1", 0))
    port = sk.getsockname()[1]
    sk.close()
    mctx = mp.get_context("spawn")
    q = mctx.Queue()
    procs = [mctx.Process(target=_host_worker, args=(r, world, port, w, h, band_rows, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = q.get(timeout=300)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert res is True
