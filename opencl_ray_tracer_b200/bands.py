"""Interleaved row-band sharding of one frame over the ranks of a torch.distributed job (one
process per GPU) -- host-side bookkeeping and the gather to rank 0.  (New: the reference is
single-device)

Band b = rows [b*band_rows, (b+1)*band_rows) belongs to rank b % world.  Rank r renders its bands
into a PACKED buffer (local row = (b // world)*band_rows + (y - b*band_rows)); rank 0 receives every
rank's packed rows back to back ("rank-major packed" buffer) and librtb's
rtb_assemble_bands_device de-interleaves them into the final frame.  The same arithmetic lives in
rtb_band_local_rows / raycast_grid_kernel / assemble_bands_kernel (csrc/).
"""
import numpy as np


def local_rows(height, band_rows, rank, world):
    """Rows rank `rank` owns (mirror of rtb_band_local_rows)."""
    n_bands = (height + band_rows - 1) // band_rows
    return sum(min(band_rows, height - b * band_rows) for b in range(rank, n_bands, world))


def global_rows_of_rank(height, band_rows, rank, world):
    """Global row index of each packed local row of `rank`, in packed order."""
    n_bands = (height + band_rows - 1) // band_rows
    out = []
    for b in range(rank, n_bands, world):
        out.extend(range(b * band_rows, min(height, (b + 1) * band_rows)))
    return np.asarray(out, dtype=np.int64)


def packed_to_frame_rows(height, band_rows, world):
    """For the rank-major packed buffer: frame_row[i] of packed row i (a permutation of 0..height-1)."""
    return np.concatenate([global_rows_of_rank(height, band_rows, r, world) for r in range(world)])


def gather_packed(dist, my_slice, packed, rows_of, rank, world):
    """Gathers every rank's packed rows into rank 0's `packed` tensor (rank 0's own rows are expected
    to be rendered in place at packed[:rows_of[0]]).  Works for CUDA tensors over NCCL and CPU
    tensors over gloo.  Grouped point-to-point: ranks own different row counts (ragged)."""
    if world == 1:
        return
    ops = []
    if rank == 0:
        off = rows_of[0]
        for r in range(1, world):
            if rows_of[r]:
                ops.append(dist.P2POp(dist.irecv, packed[off:off + rows_of[r]], r))
            off += rows_of[r]
    elif rows_of[rank]:
        ops.append(dist.P2POp(dist.isend, my_slice[:rows_of[rank]], 0))
    if ops:
        for req in dist.batch_isend_irecv(ops):
            req.wait()


class PeerFrame:
    """Rank 0's whole rgba8 frame, mapped into every rank (CUDA IPC through librtb), so that each
    rank's grid kernel stores its row bands straight into it over NVLink / NVSwitch: no gather, no
    de-interleave pass.  `dist` is torch.distributed (NCCL in production, gloo in the tests) and is
    used for the 64-byte handle broadcast and for the per-frame "every rank has finished" token."""

    def __init__(self, dist, ctx, width, height, rank, world, device=None, n_frames=1):
        import torch
        self.dist, self.ctx, self.rank, self.world = dist, ctx, rank, world
        self.width, self.height, self.n_frames = width, height, n_frames
        self.frame_bytes = 4 * width * height
        self.nbytes = self.frame_bytes * n_frames       # n_frames > 1: a frame sequence, back to back
        self.ptr, self.owner = None, rank == 0
        dev = device if device is not None else ("cuda" if dist.get_backend() == "nccl" else "cpu")
        ok, hb = 1, torch.zeros(64, dtype=torch.uint8)
        if rank == 0:
            try:
                self.ptr, handle = ctx.peer_frame_create(self.nbytes)
                hb = torch.tensor(list(handle), dtype=torch.uint8)
            except Exception:
                ok = 0
        hb = hb.to(dev)
        dist.broadcast(hb, 0)
        if rank != 0:
            try:
                self.ptr = ctx.peer_frame_open(bytes(hb.cpu().numpy().tobytes()))
            except Exception:
                ok = 0
        flag = torch.tensor([ok], dtype=torch.int32).to(dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        self.ok = bool(int(flag.item()))
        self.token = torch.zeros(1, dtype=torch.int32).to(dev)
        # signal word of the allocation (rank 0's memory): every rank's launch adds 1 when its bands have landed
        self.signal_ptr = (self.ptr + ctx.lib.rtb_peer_frame_signal_offset(self.nbytes)) if self.ptr else None
        self.signals_expected = 0
        if not self.ok:
            self.close()

    def render(self, camera, band_rows, stream=0, frame=0, signal=False):
        """This rank's bands into frame `frame` of the shared buffer (stream ordered on NCCL; the
        caller synchronises before a gloo token).  signal=True: the launch's last CTA bumps rank 0's
        signal word (see wait_all)."""
        self.ctx.render_bands_frame(self.width, self.height, camera, band_rows, self.rank, self.world,
                                    self.ptr + frame * self.frame_bytes, stream, self.signal_ptr if signal else 0)
        if signal:
            self.signals_expected += self.world

    def wait_all(self, stream=0):
        """Rank 0: `stream` waits until every rank's launches signalled so far have landed (a stream
        memory operation on rank 0's own signal word; no collective).  Raises RtbError
        (RTB_ERR_UNSUPPORTED) where stream memory operations are unavailable."""
        if self.rank == 0:
            self.ctx.stream_wait_signal(self.signal_ptr, self.signals_expected, stream)

    def fence(self):
        """Collective fence (all ranks): also what releases a frame slot for reuse."""
        self.dist.all_reduce(self.token)

    def as_tensor(self):
        """torch uint8 view [H, W, 4] of the frame (rank 0)."""
        import torch

        class _Raw:
            pass
        raw = _Raw()
        shape = (self.height, self.width, 4) if self.n_frames == 1 else (self.n_frames, self.height, self.width, 4)
        raw.__cuda_array_interface__ = {"shape": shape, "typestr": "|u1",
                                        "data": (int(self.ptr), False), "version": 3, "strides": None}
        return torch.as_tensor(raw, device="cuda")

    def close(self):
        if self.ptr:
            try:
                self.ctx.peer_frame_close(self.ptr, self.owner)
            finally:
                self.ptr = None


class SharedHostFrame:
    """One rgba8 frame in POSIX shared memory, mapped by every rank's process and page-locked in each
    (rtb_host_register), so that rtb_render_bands_host lets every GPU deliver its row bands over its
    OWN PCIe link straight into the frame rank 0 reads: the device-to-host leg of a multi-GPU frame
    scales with the number of GPUs and needs no gather.  `dist` carries the segment's name."""

    def __init__(self, dist, width, height, rank, world, register=True):
        from multiprocessing import shared_memory
        self.rank, self.world, self.width, self.height = rank, world, width, height
        nbytes = 4 * width * height
        names = [None]
        if rank == 0:
            self.shm = shared_memory.SharedMemory(create=True, size=nbytes)
            names = [self.shm.name]
        if world > 1:
            dist.broadcast_object_list(names, 0)
        if rank != 0:
            self.shm = shared_memory.SharedMemory(name=names[0])
            try:  # the creator unlinks; attached processes must not (Python < 3.13 tracks them too)
                from multiprocessing import resource_tracker
                resource_tracker.unregister(self.shm._name, "shared_memory")
            except Exception:
                pass
        self.frame = np.frombuffer(self.shm.buf, dtype=np.uint8, count=nbytes).reshape(height, width, 4)
        self.registered = False
        if register:
            from . import capi
            try:
                capi.host_register(self.frame)
                self.registered = True
            except capi.RtbError:
                pass    # not page-locked: the copies still work (staged by the driver), only slower

    def render(self, ctx, camera, band_rows):
        """This rank's bands into the shared frame (blocking); returns the kernel's device ms."""
        return ctx.render_bands_host(self.width, self.height, camera, band_rows, self.rank, self.world, self.frame)

    def close(self):
        if self.registered:
            from . import capi
            capi.host_unregister(self.frame)
            self.registered = False
        self.frame = None
        try:
            self.shm.close()
            if self.rank == 0:
                self.shm.unlink()
        except Exception:
            pass
