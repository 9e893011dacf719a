"""rtb_group_* (one process per GPU, one C call per frame and rank, no collective): every rank's kernel stores its
row bands into a slot of rank 0's frame ring (CUDA IPC: NVLink peer memory on a multi-GPU box), per-rank done words
fence the frame on rank 0's stream, per-rank credit words hand a slot back.  The ranks here are processes on
cuda:(rank % device_count): one B200 at round end (IPC between processes works on one device), distinct GPUs where
the box has them.  Nothing in the frame loop uses torch.distributed.

Checked: every frame of a sequence equals the single-GPU render although (a) the ranks enqueue the whole sequence
ahead of rank 0's consumer, (b) one rank lags (the advisor's case: with ONE summed signal word a fast rank's surplus
covered a slow rank's missing bands; with a word per rank it cannot), (c) there are fewer slots than frames in
flight, so a fast rank must wait for its credit before it overwrites a slot rank 0 is still reading."""
import os
import time

import numpy as np
import pytest

from parity import ROOT  # noqa: F401  (puts the repo root on sys.path, also for the spawned ranks)

pytestmark = pytest.mark.gpu


def _worker(rank, world, name, w, h, band_rows, slots, mode, lag_rank, q, env=None):
    os.environ.update(env or {})          # read by librtb.so when the first frame is enqueued
    import torch
    from opencl_ray_tracer_b200 import capi, host_scene as hs
    dev = rank % torch.cuda.device_count()
    torch.cuda.set_device(dev)
    ctx = capi.Context(dev)
    try:
        ctx.set_option(capi.RTB_OPT_TILE_CULL, mode)
        s = hs.random_scene(w, h, 50, 45, 31)
        ctx.set_scene(*s.as_tuple())
        g = capi.Group(ctx, name, rank, world, w, h, band_rows, slots)
        cams = [hs.orbit_camera(w, h, 3 * k) for k in range(7)]
        stream = torch.cuda.Stream()
        keep = torch.zeros((len(cams), h, w, 4), dtype=torch.uint8, device="cuda") if rank == 0 else None
        with torch.cuda.stream(stream):
            for f, cam in enumerate(cams):
                if rank == lag_rank and f in (1, 4):
                    time.sleep(0.3)                      # this rank falls behind; the others have enqueued everything
                if rank == 0 and f == 2:
                    time.sleep(0.3)                      # the consumer stalls: slots fill up, credits are withheld
                ptr = g.render(cam, stream.cuda_stream)
                if rank == 0:
                    keep[f].copy_(g.frame_tensor(ptr), non_blocking=True)   # consume the frame on the same stream
        stream.synchronize()
        g.barrier()
        ok = True
        if rank == 0:
            got = keep.cpu().numpy()
            for f, cam in enumerate(cams):
                whole = ctx.render(w, h, cam, want=("rgba8",))["rgba8"]
                if not np.array_equal(got[f], whole):
                    ok = False
                    print("device frame %d differs on %d bytes" % (f, int((got[f] != whole).sum())), flush=True)
        # host frames: every GPU delivers its bands into ONE shared, page-locked host frame
        for f, cam in enumerate(cams[:3]):
            if rank == lag_rank and f == 1:
                time.sleep(0.2)
            frame, kms = g.render_host(cam)
            if rank == 0:
                whole = ctx.render(w, h, cam, want=("rgba8",))["rgba8"]
                ok = ok and frame is not None and bool(np.array_equal(frame, whole)) and kms >= 0
            else:
                ok = ok and frame is None
        g.barrier()
        if rank == 0:
            q.put(ok)
        g.close()
    except Exception as e:  # pragma: no cover
        q.put("rank %d: %r" % (rank, e))
        raise
    finally:
        ctx.close()


# the fallbacks of the frame fence: the launch's last CTA bumps the done word (instead of a stream write behind the
# launch), and rank 0 waits on the done words one call at a time (instead of one batched stream operation)
FALLBACKS = {"RTB_GROUP_STREAM_SIGNAL": "0", "RTB_GROUP_BATCH_WAIT": "0"}


@pytest.mark.parametrize("world,w,h,band_rows,slots,mode,lag_rank,env", [
    (2, 640, 480, 16, 2, 0, 1, None),     # dense kernel, 2 slots
    (3, 333, 203, 4, 2, 0, 2, None),      # ragged frame, 4-row bands, rank 2 lags (world 3: the advisor's case)
    (3, 640, 360, 16, 3, 3, 1, None),     # the library's default mode (auto -> binned), 3 slots
    (4, 256, 64, 16, 1, 0, 3, None),      # one slot: no run-ahead at all; one band per rank
    (3, 333, 203, 4, 2, 0, 2, FALLBACKS),
    (4, 640, 360, 16, 3, 3, 1, FALLBACKS),
])
def test_group_frames_equal_single_gpu_frames(built_libs, world, w, h, band_rows, slots, mode, lag_rank, env):
    import torch.multiprocessing as mp
    mctx = mp.get_context("spawn")
    q = mctx.Queue()
    name = "rtbtest_%d_%d" % (os.getpid(), int(time.time() * 1e3) % 1000000)
    procs = [mctx.Process(target=_worker, args=(r, world, name, w, h, band_rows, slots, mode, lag_rank, q, env)) for r in range(world)]
    for p in procs:
        p.start()
    res = q.get(timeout=300)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert res is True, res


def test_group_argument_errors(gpu_ctx):
    from opencl_ray_tracer_b200 import capi
    with pytest.raises(capi.RtbError) as e:
        capi.Group(gpu_ctx, "x", 2, 2, 64, 48)        # rank out of range
    assert e.value.status == 1
    with pytest.raises(capi.RtbError):
        capi.Group(gpu_ctx, "x", 0, 1, 64, 48, band_rows=6)
    g = capi.Group(gpu_ctx, "rtbtest_solo_%d" % os.getpid(), 0, 1, 64, 48)   # a group of one works like a context
    try:
        from opencl_ray_tracer_b200 import host_scene as hs
        gpu_ctx.set_scene(*hs.random_scene(64, 48, 3, 3, 1).as_tuple())
        frame, _ = g.render_host(capi.REFERENCE_CAMERA)
        assert np.array_equal(frame, gpu_ctx.render(64, 48, want=("rgba8",))["rgba8"])
        g.barrier()
    finally:
        g.close()


def test_multi_device_context(built_libs):
    """rtb_init(n_devices > 1): ONE blocking rtb_render per frame shards over the GPUs of this box (needs >= 2)."""
    import torch
    from opencl_ray_tracer_b200 import capi, host_scene as hs
    n = min(torch.cuda.device_count(), 4)
    if n < 2:
        with pytest.raises(capi.RtbError) as e:
            capi.Context(devices=[0, 1])
        assert e.value.status == 2                      # RTB_ERR_NO_DEVICE: loud, no fallback to one GPU
        pytest.skip("one GPU on this box")
    w, h = 1920, 1080
    s = hs.random_scene(w, h, 64, 64, 2)
    with capi.Context(0) as one:
        one.set_scene(*s.as_tuple())
        ref = one.render(w, h)
    with capi.Context(devices=list(range(n))) as multi:
        multi.set_scene(*s.as_tuple())
        for mode in (capi.TILE_CULL_DENSE, capi.TILE_CULL_AUTO):
            multi.set_option(capi.RTB_OPT_TILE_CULL, mode)
            got = multi.render(w, h)
            for k in ("rgba8", "prim_id", "rgba_i32"):
                assert np.array_equal(got[k], ref[k]), (mode, k)
        assert multi.launch_count() >= 2 * n
        cams = [hs.orbit_camera(w, h, 5 * k) for k in range(4)]
        out = torch.zeros((len(cams), h, w, 4), dtype=torch.uint8, device="cuda:0")
        multi.render_sequence_device(w, h, np.stack([np.concatenate(c) for c in cams]), out.data_ptr())
        seq = out.cpu().numpy()
    with capi.Context(0) as one:
        one.set_scene(*s.as_tuple())
        for k, c in enumerate(cams):
            assert np.array_equal(seq[k], one.render(w, h, c, want=("rgba8",))["rgba8"]), k
