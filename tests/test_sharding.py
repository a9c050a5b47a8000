"""Host-side sharding logic, covered on CPU: world_size-2 and -3 gloo groups
exercise the same gather_rows / reduce_samples code the NCCL path runs, on synthetic frames
whose rows / sample ranges are recognisable."""
import os
import socket
import sys

import numpy as np
import pytest

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from opengl_raytracing_b200 import sharding  # noqa: E402


@pytest.mark.parametrize("height,world", [(1080, 1), (1080, 2), (1080, 8), (4320, 8), (7, 4), (4, 8), (2160, 3)])
def test_row_ranges_partition_the_frame(height, world):
    rows = []
    for r in range(world):
        a, b = sharding.row_range(height, r, world)
        assert 0 <= a <= b <= height
        assert a % 4 == 0 or a == height              # no 8x4 warp tile straddles two ranks
        rows += list(range(a, b))
    assert rows == list(range(height))


@pytest.mark.parametrize("n,world", [(64, 1), (64, 2), (64, 8), (5, 8), (7, 3)])
def test_sample_ranges_partition_the_samples(n, world):
    got = []
    for r in range(world):
        first, count = sharding.sample_range(n, r, world)
        got += list(range(first, first + count))
    assert got == list(range(n))


def _free_port() -> int:
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank: int, world: int, port: int, height: int, out_dir: str):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        W = 16
        # --- row tiles: each rank fills only its own rows with (row index, rank)
        buf = torch.full((height, W, 4), -1.0)
        a, b = sharding.row_range(height, rank, world)
        for y in range(a, b):
            buf[y, :, 0] = float(y)
            buf[y, :, 1] = float(rank)
        sharding.gather_rows(buf, height, world, dst=0)
        if rank == 0:
            np.save(os.path.join(out_dir, "gathered.npy"), buf.numpy())
        # --- sample ranges: rank g contributes sum over its frame_count range of (s + 1)
        first, count = sharding.sample_range(10, rank, world)
        part = torch.zeros(3, 5, 4)
        for s in range(first, first + count):
            part[..., :3] += float(s + 1)
        part[..., 3] = 123.0                                  # alpha is restored to 1 by the reduce
        res = sharding.reduce_samples(part, 10, dst=0)
        if rank == 0:
            np.save(os.path.join(out_dir, "reduced.npy"), res.numpy())
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,height", [(2, 40), (3, 22)])
def test_gather_and_reduce_over_gloo(tmp_path, world, height):
    import torch.multiprocessing as mp
    port = _free_port()
    mp.spawn(_worker, args=(world, port, height, str(tmp_path)), nprocs=world, join=True)
    g = np.load(tmp_path / "gathered.npy")
    assert np.array_equal(g[:, 0, 0], np.arange(height, dtype=np.float32))       # every row arrived, in place
    for r in range(world):
        a, b = sharding.row_range(height, r, world)
        assert (g[a:b, :, 1] == r).all()
    red = np.load(tmp_path / "reduced.npy")
    assert np.allclose(red[..., :3], sum(range(1, 11)) / 10.0) and (red[..., 3] == 1.0).all()
