"""N > 1 host logic on CPU: world_size-2 (and 3) process groups over gloo.  Each rank fabricates the output its shard
would produce (a deterministic function of the pixel index, laid out exactly as rtc_render lays out a shard: whole
full-width tiles in ascending order, partial last band padded) and the two assembly paths of sharding.py must both
rebuild the full frame: the shared host image (what bench.py uses) and the torch.distributed gather."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import raytracer_challenge_b200 as R
from raytracer_challenge_b200 import sharding


def fake_pixel(idx):
    idx = idx.astype(np.float64)
    return np.stack([idx, idx * 0.5 + 1.0, -idx], axis=-1)


def fake_shard_output(w, h, rank, world):
    idx = R.shard_pixels(w, h, rank, world, (w, sharding.TILE_H))  # slot -> linear pixel index or 0xffffffff
    out = fake_pixel(idx)
    out[idx == 0xFFFFFFFF] = np.nan  # padding slots of the partial last band must never reach the frame
    return out.reshape(-1, w, 3)


def worker(rank, world, port, w, h, shm_name, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        local = fake_shard_output(w, h, rank, world)
        # rows owned by this rank, as the band arithmetic says
        rows = sharding.band_rows(h, rank, world)
        assert local.shape[0] >= len(rows)
        # path 1: shared host frame
        frame = sharding.SharedFrame(h, w, name=shm_name, create=False)
        sharding.place_bands(frame.array, local, rank, world)
        frame.frame_barrier(rank, world)  # the per-frame barrier of the end-to-end bench: counters in the shared segment
        if rank == 0:
            want = fake_pixel(np.arange(w * h)).reshape(h, w, 3)
            q.put(("shared", bool(np.array_equal(frame.array, want))))
        dist.barrier()
        frame.close()
        # path 2: gather to rank 0
        t = torch.from_numpy(np.nan_to_num(local, nan=-7.0))
        full = sharding.gather_bands(t, h, w, rank, world)
        if rank == 0:
            want = fake_pixel(np.arange(w * h)).reshape(h, w, 3)
            q.put(("gather", bool(np.array_equal(full.numpy(), want))))
        else:
            assert full is None
        # opts every rank would pass to rtc_render
        o = sharding.band_opts(w, rank, world)
        assert (o.shard_index, o.shard_count, o.tile_w, o.tile_h) == (rank, world, w, sharding.TILE_H)
        assert o.flags & R.FLAG_SCATTER_ROWS
    finally:
        dist.destroy_process_group()


def free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("world,w,h", [(2, 64, 36), (2, 33, 19), (3, 40, 10)])
def test_two_rank_assembly(built, world, w, h):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    frame = sharding.SharedFrame(h, w, create=True)
    frame.array[:] = -1.0
    try:
        procs = [ctx.Process(target=worker, args=(r, world, PORT[0], w, h, frame.name, q)) for r in range(world)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(timeout=120)
            assert p.exitcode == 0
        results = dict(q.get(timeout=10) for _ in range(2))
        assert results == {"shared": True, "gather": True}
    finally:
        frame.close()


PORT = [0]


@pytest.fixture(autouse=True)
def _port():
    PORT[0] = free_port()


def test_band_rows_partition():
    for h, world in [(36, 2), (19, 2), (10, 3), (2160, 8), (5, 8)]:
        seen = np.zeros(h, dtype=int)
        for r in range(world):
            seen[sharding.band_rows(h, r, world)] += 1
        assert (seen == 1).all()
