"""The device-consumer form of the multi-GPU gather on real GPUs: every rank renders its bands of the frame on its own GPU
(rtc_render through the C ABI), the bands are gathered to rank 0 over NCCL (NVLink / NVSwitch) with sharding.gather_bands and
interleaved there; the result must equal the frame one GPU renders alone, bit for bit.  Needs >= 2 GPUs (skipped on a
one-GPU box; run by tools/gpu.sh multi-box calls)."""
import os
import socket

import numpy as np
import pytest
import torch

import raytracer_challenge_b200 as R
from raytracer_challenge_b200 import sharding

pytestmark = pytest.mark.gpu


def worker(rank, world, port, name, w, h, q):
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        with R.Session(name, w, h, device=rank) as s:
            part, st = s.render(R.Opts(shard=(rank, world), tile=(w, sharding.TILE_H)))  # this rank's bands: [n_bands * TILE_H, w, 3]
            whole = s.render(R.Opts())[0] if rank == 0 else None
        local = torch.from_numpy(np.nan_to_num(part.reshape(-1, w, 3), nan=0.0)).cuda()
        full = sharding.gather_bands(local, h, w, rank, world)
        if rank == 0:
            q.put(bool(np.array_equal(full.cpu().numpy(), whole)))
        else:
            assert full is None
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 4])
def test_bands_gathered_over_nccl_equal_the_whole_frame(built, world):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    import torch.multiprocessing as mp

    with socket.socket() as sock:
        sock.bind(("127.0.0.1", 0))
        port = sock.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=worker, args=(r, world, port, "teapot_low", 96, 54, q)) for r in range(world)]
    for p in procs:
        p.start()
    ok = q.get(timeout=300)
    for p in procs:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert ok
