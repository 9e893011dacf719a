"""N > 1 host logic on CPU: band bookkeeping (python mirror vs the C ABI's rtb_band_local_rows) and
the gather to rank 0 over torch.distributed with the gloo backend, world_size 2 and 3."""
import os
import socket

import numpy as np
import pytest

from parity import ROOT


@pytest.mark.parametrize("h,band_rows,world", [(2160, 16, 8), (1083, 16, 4), (203, 4, 3), (10, 16, 2), (4320, 16, 1)])
def test_band_bookkeeping_matches_c_abi(built_libs, h, band_rows, world):
    from opencl_ray_tracer_b200 import bands, capi
    lib = capi.load_library()
    rows = [bands.local_rows(h, band_rows, r, world) for r in range(world)]
    assert rows == [lib.rtb_band_local_rows(h, band_rows, r, world) for r in range(world)]
    perm = bands.packed_to_frame_rows(h, band_rows, world)
    assert sorted(perm.tolist()) == list(range(h))            # every frame row exactly once
    for r in range(world):
        g = bands.global_rows_of_rank(h, band_rows, r, world)
        assert len(g) == rows[r] and all((y // band_rows) % world == r for y in g)


def _worker(rank, world, port, h, w, band_rows, q):
    import torch
    import torch.distributed as dist
    from opencl_ray_tracer_b200 import bands
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        frame = torch.arange(h * w * 4, dtype=torch.int64).reshape(h, w, 4).remainder(251).to(torch.uint8)  # the "image"
        rows_of = [bands.local_rows(h, band_rows, r, world) for r in range(world)]
        mine = torch.from_numpy(bands.global_rows_of_rank(h, band_rows, rank, world))
        if rank == 0:
            packed = torch.zeros((h, w, 4), dtype=torch.uint8)
            my_slice = packed[:rows_of[0]]
        else:
            packed, my_slice = None, torch.zeros((max(rows_of[rank], 1), w, 4), dtype=torch.uint8)
        if rows_of[rank]:
            my_slice[:rows_of[rank]] = frame[mine]          # what the band kernel writes for this rank
        bands.gather_packed(dist, my_slice, packed, rows_of, rank, world)
        if rank == 0:
            out = torch.zeros_like(frame)
            out[torch.from_numpy(bands.packed_to_frame_rows(h, band_rows, world))] = packed  # assemble kernel's job
            q.put(bool(torch.equal(out, frame)))
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,h,band_rows", [(2, 203, 16), (3, 100, 4), (2, 10, 16)])
def test_gather_over_gloo(world, h, band_rows):
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, h, 37, band_rows, q)) for r in range(world)]
    for p in procs:
        p.start()
    ok = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert ok


def _shared_worker(rank, world, port, h, w, band_rows, q):
    import torch.distributed as dist
    from opencl_ray_tracer_b200 import bands
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        # no device here: the segment is created / attached / unlinked, not page-locked
        shf = bands.SharedHostFrame(dist, w, h, rank, world, register=False)
        want = (np.arange(h * w * 4, dtype=np.int64).reshape(h, w, 4) % 251).astype(np.uint8)
        mine = bands.global_rows_of_rank(h, band_rows, rank, world)
        shf.frame[mine] = want[mine]                      # what rtb_render_bands_host delivers for this rank
        dist.barrier()
        if rank == 0:
            q.put(bool(np.array_equal(shf.frame, want)))
        dist.barrier()
        shf.close()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,h,band_rows", [(2, 203, 16), (3, 100, 4)])
def test_shared_host_frame_over_gloo(world, h, band_rows):
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_shared_worker, args=(r, world, port, h, 37, band_rows, q)) for r in range(world)]
    for p in procs:
        p.start()
    ok = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert ok
