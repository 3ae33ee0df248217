"""Multi-GPU plumbing: chains are independent, so the path shards with NO
data-path collective -- rank r runs chains [r*C_local, (r+1)*C_local) with the Philox
counter keyed by the GLOBAL chain index (results do not depend on the GPU count) -- and
the only exchange is ONE all-reduce (sum) of the posterior statistics at the end of a
run: mixtures {count, S(theta), S(theta theta^T), 100x100 histogram, accepted count};
images {count, S(x), S(x^2)} pooled over chains.  ``torch.distributed`` is the plumbing
(NCCL on the GPU box, gloo in the CPU tests); payloads are KB..MB, latency-bound.
"""
from __future__ import annotations

import os


def world():
    """(rank, world_size, local_rank) from the torchrun environment."""
    return (int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")),
            int(os.environ.get("LOCAL_RANK", "0")))


def shard(total, rank, world_size):
    """Contiguous split of ``total`` chains: (offset, count) of this rank; the first
    ``total % world_size`` ranks take one extra chain."""
    base, rem = divmod(int(total), int(world_size))
    count = base + (1 if rank < rem else 0)
    offset = rank * base + min(rank, rem)
    return offset, count


def _dist():
    import torch.distributed as dist
    return dist if (dist.is_available() and dist.is_initialized()) else None


def allreduce_chain_stats_(stats, group=None):
    """In-place sum over ranks of a ``ChainStats``' moments / histogram / chain and accept
    counts.  No-op when torch.distributed is not initialised."""
    import torch
    dist = _dist()
    if dist is None or dist.get_world_size(group) == 1:
        return stats
    dist.all_reduce(stats.moments_dev, group=group)
    if stats.hist_dev is not None:
        dist.all_reduce(stats.hist_dev, group=group)
    extra = torch.zeros(2, dtype=torch.float64, device=stats.moments_dev.device)
    extra[0] = float(stats.chains)
    if stats.naccept is not None:
        extra[1] = stats.naccept.double().sum().to(extra.device)
    dist.all_reduce(extra, group=group)
    stats.chains = int(round(float(extra[0].item())))
    stats.global_accepted = float(extra[1].item()) if stats.naccept is not None else None
    stats._host = None
    return stats


def allreduce_image_sums(count, s1, s2=None, group=None):
    """Sum over ranks of the pooled image statistics (``ImageStats.pooled_sums()``); returns
    (count, mean image, second-moment image or None) as tensors on the input device."""
    import torch
    dist = _dist()
    n = torch.tensor([float(count)], dtype=torch.float64, device=s1.device)
    if dist is not None and dist.get_world_size(group) > 1:
        dist.all_reduce(n, group=group)
        dist.all_reduce(s1, group=group)
        if s2 is not None:
            dist.all_reduce(s2, group=group)
    tot = float(n.item())
    return tot, s1 / tot, (None if s2 is None else s2 / tot)
