"""N > 1 host logic on CPU: world_size-2 gloo processes shard a batch, each 'processes' its range
(a NumPy stand-in for the per-image checksum kernel), and the gathered per-image statistics
equal the single-process result for even and uneven splits."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from image_denoising_b200 import shard


def _checksum(img):
    flat = img.reshape(-1).astype(np.uint64)
    return int((flat * (1 + (np.arange(flat.size, dtype=np.uint64) % 251))).sum() & 0x7FFFFFFFFFFFFFFF)


def _worker(rank, world, port, n, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        start, cnt = shard.my_range(n)
        rng = np.random.RandomState(0)
        batch = rng.randint(0, 256, (n, 8, 12, 3)).astype(np.uint8)       # same data on every rank
        local = torch.tensor([[_checksum(batch[start + i]), start + i] for i in range(cnt)], dtype=torch.int64)
        local = local.reshape(cnt, 2)
        full = shard.gather_per_image(local, n)
        if rank == 0:
            q.put(full.numpy())
    finally:
        dist.destroy_process_group()


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("n", [8, 7, 1])
def test_two_rank_gather_equals_single_process(n):
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, n, q)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    got = q.get()
    rng = np.random.RandomState(0)
    batch = rng.randint(0, 256, (n, 8, 12, 3)).astype(np.uint8)
    want = np.array([[_checksum(batch[i]), i] for i in range(n)], dtype=np.int64)
    assert np.array_equal(got, want)


def test_partition_properties():
    for n in (0, 1, 7, 512, 4096):
        for world in (1, 2, 4, 8):
            parts = shard.partition(n, world)
            assert len(parts) == world and sum(c for _, c in parts) == n
            assert all(parts[i][0] + parts[i][1] == parts[i + 1][0] for i in range(world - 1))
            assert max(c for _, c in parts) - min(c for _, c in parts) <= 1
