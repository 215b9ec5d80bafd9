"""Multi-GPU sharding of a batch: chunks are independent, so every rank takes a
contiguous range of the batch, balanced by BYTES, and runs the single-GPU path on it.  There is no
collective on the data path; the only exchange is the size/offset gather below (8 B per chunk),
needed when every rank must know where its compressed range lands in one global output
(e.g. to assemble a single HLIF container from N GPUs).

Pure host logic over torch.distributed: runs with NCCL on the GPU box and with gloo in the CPU tests.
"""
import numpy as np
import torch
import torch.distributed as dist


def shard_bounds(sizes, world):
    """Boundaries b[0..world] of contiguous chunk ranges with near-equal byte totals:
    rank r owns chunks [b[r], b[r+1]).  Deterministic, identical on every rank."""
    sizes = np.asarray(sizes, dtype=np.int64)
    n = len(sizes)
    if world <= 0:
        raise ValueError("world must be positive")
    csum = np.concatenate([[0], np.cumsum(sizes)])
    total = int(csum[-1])
    bounds = [0]
    for r in range(1, world):
        target = total * r / world
        # first boundary whose prefix reaches the target, never before the previous boundary
        b = int(np.searchsorted(csum, target, side="left"))
        if b > 0 and abs(csum[b - 1] - target) <= abs(csum[min(b, n)] - target):
            b -= 1
        bounds.append(min(max(b, bounds[-1]), n))
    bounds.append(n)
    return np.asarray(bounds, dtype=np.int64)


def my_range(sizes, rank, world):
    b = shard_bounds(sizes, world)
    return int(b[rank]), int(b[rank + 1])


def gather_sizes(local_sizes, group=None):
    """All ranks contribute the (compressed) sizes of their chunk range; returns
    (global_sizes int64[n_total], global_offsets int64[n_total], my_first_chunk, my_base_offset).
    One all_gather of the counts + one all_gather of the padded size arrays."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    local = torch.as_tensor(local_sizes, dtype=torch.int64)
    dev = local.device
    count = torch.tensor([local.numel()], dtype=torch.int64, device=dev)
    counts = [torch.zeros_like(count) for _ in range(world)]
    dist.all_gather(counts, count, group=group)
    counts = [int(c.item()) for c in counts]
    width = max(max(counts), 1)
    padded = torch.zeros(width, dtype=torch.int64, device=dev)
    padded[: local.numel()] = local
    parts = [torch.zeros_like(padded) for _ in range(world)]
    dist.all_gather(parts, padded, group=group)
    sizes = torch.cat([p[:c] for p, c in zip(parts, counts)]).cpu()
    offsets = torch.cumsum(sizes, 0) - sizes
    first = sum(counts[:rank])
    base = int(offsets[first].item()) if first < sizes.numel() else int(sizes.sum().item())
    return sizes.numpy(), offsets.numpy(), first, base
