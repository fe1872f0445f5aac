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


def bucket_by_bands(patches):
    """Mixed-sensor batches (BASELINE config 3): stacked tensors need equal band counts, and zero-padding 4 -> 8 bands
    is not equivalent (biases and pad-1 convolutions leak into padded bands), so a joint batch is split
    into one bucket per band count.  `patches`: sequence of (PAN [1,H,W], MS [N,H,W], prompt).  Returns
    {N: (indices, PAN [b,1,H,W], MS [b,N,H,W], prompts)}; per-sample prompts inside a bucket are free (the text
    modulation is a per-sample scale table)."""
    buckets = {}
    for i, (pan, ms, prompt) in enumerate(patches):
        buckets.setdefault(int(ms.shape[0]), []).append(i)
    out = {}
    for n, idx in sorted(buckets.items()):
        out[n] = (idx, torch.stack([patches[i][0] for i in idx]), torch.stack([patches[i][1] for i in idx]),
                  [patches[i][2] for i in idx])
    return out


@torch.no_grad()
def sample_buckets(diffusion, buckets, steps=20, group=None, batch=64):
    """The sharded core of sample_mixed.  buckets: {bands: (PAN [b,1,H,W], MS [b,bands,H,W], prompts (b names), x_T [b,bands,H,W])},
    every rank holding the same full buckets on the host (pinned: the local shard is copied asynchronously) or on the device.
    Each bucket is split contiguously over the ranks, the local shard sampled in chunks of at most `batch` patches, the results
    all-gathered (one collective per bucket).  Returns {bands: fused images [b,bands,H,W]} on the device."""
    on = dist.is_available() and dist.is_initialized()
    rank, world = (dist.get_rank(group), dist.get_world_size(group)) if on else (0, 1)
    betas = getattr(diffusion, "betas", None)          # the sampler's device; test doubles without a schedule stay where the data is
    out = {}
    for n in sorted(buckets):
        pan, ms, prompts, x_T = buckets[n]
        dev = betas.device if betas is not None else ms.device
        lo, hi = shard_range(ms.shape[0], rank, world)
        outs = []
        for s in range(lo, hi, batch):
            e = min(s + batch, hi)
            p, m = pan[s:e].to(dev, non_blocking=True), ms[s:e].to(dev, non_blocking=True)
            noise = torch.randn_like(m) if x_T is None else x_T[s:e].to(dev, non_blocking=True)
            outs.append(diffusion.sample_by_dpmsolver({"PAN": p, "MS": m, "Res": noise}, list(prompts[s:e]), steps=steps, x_T=noise))
        local = torch.cat(outs) if outs else torch.zeros((0,) + tuple(ms.shape[1:]), device=dev)
        out[n] = gather_tiles(local, ms.shape[0], group) if world > 1 else local
    return out


@torch.no_grad()
def sample_mixed(diffusion, patches, steps=20, x_T=None, group=None, batch=64):
    """Sample a joint WV3/QB/GF2 batch (BASELINE config 3): bucket by band count, shard every bucket contiguously over the ranks
    (equal FLOPs per rank: an 8-band patch costs twice a 4-band one, and every rank gets the same share of each bucket), run the
    local shard in chunks of at most `batch` patches with per-sample prompts (the module keeps one engine per band count, so
    alternating buckets re-plans nothing), one all-gather per bucket.  Returns the fused images in input order (list of
    [N_i, H, W])."""
    by_bands = bucket_by_bands(patches)
    buckets = {n: (pan, ms, prompts, None if x_T is None else torch.stack([x_T[i] for i in idx]))
               for n, (idx, pan, ms, prompts) in by_bands.items()}
    fused = sample_buckets(diffusion, buckets, steps=steps, group=group, batch=batch)
    result = [None] * len(patches)
    for n, (idx, _, _, _) in by_bands.items():
        for j, i in enumerate(idx):
            result[i] = fused[n][j]
    return result
