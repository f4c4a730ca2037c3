"""Batch sharding of clips over the GPUs of one box: clips are independent along the
batch dimension (per-frame spatial attention batches over (n t), temporal attention over (n s), LayerNorm
only), so inference needs no data-path collective.  One process per GPU, as `accelerate launch` does for the
reference (config/accelerate_config_8.yaml)."""
import torch


def shard_indices(n_clips, rank, world):
    """clip i goes to rank i % world (round-robin keeps shards within one clip of each other)."""
    if not (0 <= rank < world):
        raise ValueError("rank %d outside world of %d" % (rank, world))
    return list(range(rank, n_clips, world))


def shard_clips(z, rank, world):
    """(n, ...) clip tensor -> this rank's clips (a copy; may be empty)."""
    idx = shard_indices(z.shape[0], rank, world)
    return z[idx] if idx else z[:0]


def gather_clips(local, n_clips, rank, world, group=None):
    """inverse of shard_clips on every rank: all-gather the per-rank results (result collection only — not
    part of the timed data path) and restore the original clip order."""
    import torch.distributed as dist
    if world == 1:
        return local
    counts = [len(shard_indices(n_clips, r, world)) for r in range(world)]
    pad = max(counts)
    buf = torch.zeros((pad,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    buf[:local.shape[0]] = local
    parts = [torch.empty_like(buf) for _ in range(world)]
    dist.all_gather(parts, buf, group=group)
    out = torch.empty((n_clips,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    for r in range(world):
        idx = shard_indices(n_clips, r, world)
        if idx:
            out[idx] = parts[r][:len(idx)]
    return out


def max_over_ranks(ms, device, world):
    """bench timing rule: a multi-GPU step takes as long as its slowest rank"""
    if world == 1:
        return float(ms)
    import torch.distributed as dist
    t = torch.tensor([float(ms)], device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)
