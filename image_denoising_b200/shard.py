"""Batch sharding across the GPUs of one box.

Every image is an independent unit: the batch is cut in contiguous ranges, one per rank, and
each rank runs `b2d_run` on its own range with `first_global_image_index = start` - the Philox
counter is keyed by the GLOBAL image index, so the result is identical for any world size.  No
collective touches pixels; the only exchange is an all_gather of per-image (checksum, sse),
16 bytes per image, after the run (NCCL over NVLink on the GPU box, gloo in the CPU tests)."""
import torch
import torch.distributed as dist


def partition(n, world):
    """Contiguous (start, count) per rank; the first n % world ranks get one extra image."""
    base, extra = divmod(int(n), int(world))
    out, start = [], 0
    for r in range(world):
        cnt = base + (1 if r < extra else 0)
        out.append((start, cnt))
        start += cnt
    return out


def my_range(n, world=None, rank=None):
    world = dist.get_world_size() if world is None else world
    rank = dist.get_rank() if rank is None else rank
    return partition(n, world)[rank]


def gather_per_image(local, n, group=None):
    """all_gather of a per-image tensor (leading dim = this rank's count) into the full
    (n, ...) tensor in global image order, on every rank.  Ranges may be uneven."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return local
    world = dist.get_world_size(group)
    parts = partition(n, world)
    width = max(c for _, c in parts)
    pad = torch.zeros((width,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    pad[:local.shape[0]] = local
    bufs = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(bufs, pad, group=group)
    return torch.cat([b[:c] for b, (_, c) in zip(bufs, parts)], 0)
