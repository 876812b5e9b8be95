"""One-process-per-GPU sharding of a frame (harness plumbing over torch.distributed; the product is the C ABI).

Pixels are independent (src/image.rs:66-74), so a frame shards without any exchange step: the window is cut into
bands of `tile_h` pixel rows (full-width tiles), band g belongs to rank g % world.  Interleaving at 4-row granularity
balances the very uneven per-pixel cost (background 1 ray, teapot up to 18).  Two ways to bring the bands together:

  * SharedFrame  — every rank maps the same page-locked host image (POSIX shared memory + cudaHostRegister) and its
                   rtc_render call (RTC_FLAG_SCATTER_ROWS) copies its bands straight to their row-major place:
                   N device->host copies over N PCIe links in parallel, no collective.  This is what bench.py times.
  * gather_bands — torch.distributed gather of the per-rank band lists to rank 0 (NCCL over NVLink for device
                   tensors, gloo for CPU tensors in the tests) followed by the same interleave; for a device consumer.
"""
from multiprocessing import shared_memory

import numpy as np

from .binding import FLAG_SCATTER_ROWS, Opts, load_rtc

TILE_H = 4


def band_opts(width, rank, world, fuel=5, spp=(1, 1), flags=0, tile_h=TILE_H):
    """rtc_opts of rank `rank`: full-width tiles of tile_h rows dealt round-robin, rows scattered into a full frame."""
    if world <= 1:
        return Opts(fuel=fuel, spp=spp, flags=flags)
    return Opts(fuel=fuel, spp=spp, shard=(rank, world), tile=(width, tile_h), flags=flags | FLAG_SCATTER_ROWS)


def band_rows(height, rank, world, tile_h=TILE_H):
    """Pixel rows owned by `rank`, in the order its shard output lists them."""
    rows = []
    n_bands = (height + tile_h - 1) // tile_h
    for g in range(rank, n_bands, world):
        rows.extend(range(g * tile_h, min((g + 1) * tile_h, height)))
    return np.asarray(rows, dtype=np.int64)


def place_bands(frame, shard_out, rank, world, tile_h=TILE_H):
    """Host-side equivalent of RTC_FLAG_SCATTER_ROWS: shard_out is the rank's plain shard output (whole tiles, i.e.
    n_bands * tile_h rows of `width` pixels, the last band possibly padded); writes its rows into frame[h, w, 3]."""
    h, w = frame.shape[0], frame.shape[1]
    rows = shard_out.reshape(-1, w, frame.shape[2])
    n_bands = (h + tile_h - 1) // tile_h
    k = 0
    for g in range(rank, n_bands, world):
        y0, y1 = g * tile_h, min((g + 1) * tile_h, h)
        frame[y0:y1] = rows[k * tile_h:k * tile_h + (y1 - y0)]
        k += 1
    return frame


def gather_bands(local, height, width, rank, world, group=None, dst=0, tile_h=TILE_H):
    """torch.distributed gather of every rank's shard output (a tensor of whole bands: [n_bands_r * tile_h, width, 3])
    to rank `dst`, interleaved there into the row-major frame.  Works for CUDA tensors (NCCL) and CPU tensors (gloo).
    Returns the [height, width, 3] frame on dst, None elsewhere."""
    import torch
    import torch.distributed as dist

    n_bands = (height + tile_h - 1) // tile_h
    max_bands = (n_bands + world - 1) // world
    padded = torch.zeros((max_bands * tile_h, width, local.shape[-1]), dtype=local.dtype, device=local.device)
    padded[: local.shape[0]] = local
    if world == 1:
        parts = [padded]
    else:
        parts = [torch.empty_like(padded) for _ in range(world)] if rank == dst else None
        dist.gather(padded, parts, dst=dst, group=group)
    if rank != dst:
        return None
    frame = torch.empty((n_bands * tile_h, width, local.shape[-1]), dtype=local.dtype, device=local.device)
    view = frame.view(n_bands, tile_h, width, local.shape[-1])
    for r in range(world):
        mine = view[r::world]
        mine.copy_(parts[r][: mine.shape[0] * tile_h].view(mine.shape[0], tile_h, width, local.shape[-1]))
    return frame[:height]


class SharedFrame:
    """A [height, width, 3] f64 host image shared by the ranks of one node (POSIX shared memory), page-locked in every
    process that calls pin() so cudaMemcpy2DAsync into it runs at full PCIe rate and asynchronously."""

    def __init__(self, height, width, name=None, create=False):
        self.shape = (height, width, 3)
        nbytes = int(np.prod(self.shape)) * 8
        flags_at = (nbytes + 4095) // 4096 * 4096  # per-rank frame counters live behind the image (frame_barrier)
        self.shm = shared_memory.SharedMemory(name=name, create=create, size=flags_at + self.MAX_RANKS * 8 if create else 0)
        self.owner = create
        self.array = np.ndarray(self.shape, dtype=np.float64, buffer=self.shm.buf)
        self.flags = np.ndarray((self.MAX_RANKS,), dtype=np.int64, buffer=self.shm.buf, offset=flags_at)
        if create:
            self.flags[:] = 0
        self.seq = 0
        self.pinned = False

    MAX_RANKS = 64

    @property
    def name(self):
        return self.shm.name

    def frame_barrier(self, rank, world, timeout_s=60.0):
        """Every rank calls this once per frame, after its bands have landed in the image (its render call has
        returned): returns when all `world` ranks have.  A counter per rank in the shared segment; each rank writes
        only its own (aligned 8-byte stores), so no atomics are needed.  A node-local stand-in for a collective
        barrier, which costs milliseconds per call here."""
        import time
        self.seq += 1
        self.flags[rank] = self.seq
        t0 = time.perf_counter()
        spins = 0
        # (plain aligned 8-byte loads / stores: relies on x86's total store order and on the rank's copies having completed
        # — its render call has returned — before its flag is stored.  After a short spin the wait backs off so that it does
        # not hold a core the other ranks' host work could use.)
        while int(self.flags[:world].min()) < self.seq:
            spins += 1
            if spins > 200:
                time.sleep(20e-6)
            if time.perf_counter() - t0 > timeout_s:
                raise TimeoutError("SharedFrame.frame_barrier: a rank did not arrive")

    @property
    def ptr(self):
        return self.array.ctypes.data

    def pin(self):
        lib = load_rtc()
        if lib.rtc_host_register(self.ptr, self.array.nbytes) != 0:
            raise RuntimeError("rtc_host_register: " + lib.rtc_last_error().decode())
        self.pinned = True

    def close(self):
        if self.pinned:
            load_rtc().rtc_host_unregister(self.ptr)
            self.pinned = False
        self.array = None
        self.flags = None
        self.shm.close()
        if self.owner:
            self.shm.unlink()
