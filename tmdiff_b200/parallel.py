"""Multi-GPU sharding of the sampling path: one process per GPU, patches split contiguously,
no per-step communication, one all-gather of the output tiles at the end.
The reference has only nn.DataParallel for training (GeneralModel/networks.py:88-91)."""
import torch
import torch.distributed as dist


def shard_range(n_items, rank, world_size):
    """Contiguous, balanced split: the first (n % world) ranks get one extra item."""
    base, extra = divmod(n_items, world_size)
    start = rank * base + min(rank, extra)
    return start, start + base + (1 if rank < extra else 0)


def gather_tiles(local, n_total, group=None):
    """All-gather ragged first-dimension shards produced by `shard_range` back into
    [n_total, ...] on every rank (NCCL over NVLink on GPUs, gloo on CPU)."""
    world = dist.get_world_size(group)
    if world == 1:
        return local
    sizes = [shard_range(n_total, r, world) for r in range(world)]
    mx = max(e - s for s, e in sizes)
    pad = local.new_zeros((mx,) + tuple(local.shape[1:]))
    pad[: local.shape[0]] = local
    out = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(out, pad, group=group)
    return torch.cat([o[: e - s] for o, (s, e) in zip(out, sizes)], dim=0)
