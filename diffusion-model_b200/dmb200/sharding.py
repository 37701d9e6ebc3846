"""Sample-sharded sampling across the GPUs of one box (BASELINE config 4).

Every sample of the path is independent end to end (per-sample GroupNorm, per-token
LayerNorm, per-sample attention), so the global batch is split into contiguous shards, one
process per GPU, with NO communication during the T-step loop; the only collective is one
all-gather of the decoded u8 images (NCCL on GPUs; the same code runs over gloo in the CPU
tests).  Noise is keyed by the GLOBAL sample index, so the gathered result does not depend
on the number of ranks.
"""
from __future__ import annotations

from typing import Callable, Optional, Tuple

import torch
import torch.distributed as dist


def shard_range(n_total: int, rank: int, world: int) -> Tuple[int, int]:
    """[begin, end) of rank's contiguous shard; the first n_total % world ranks get one extra."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError(f"bad rank/world: {rank}/{world}")
    if n_total < 0:
        raise ValueError("n_total must be >= 0")
    base, extra = divmod(n_total, world)
    begin = rank * base + min(rank, extra)
    return begin, begin + base + (1 if rank < extra else 0)


def gather_shards(local: torch.Tensor, n_total: int, group=None) -> torch.Tensor:
    """All-gather of per-rank shards (dim 0, sizes from shard_range) into the full
    (n_total, ...) tensor on every rank.  Ragged shards are padded to the largest one for
    the collective and trimmed afterwards."""
    if not (dist.is_available() and dist.is_initialized()):
        if local.shape[0] != n_total:
            raise ValueError("single process: the local shard must be the whole batch")
        return local
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    b, e = shard_range(n_total, rank, world)
    if local.shape[0] != e - b:
        raise ValueError(f"rank {rank} holds {local.shape[0]} samples, expected {e - b}")
    biggest = (n_total + world - 1) // world
    send = local.contiguous()
    if send.shape[0] != biggest:
        pad = torch.zeros((biggest - send.shape[0],) + tuple(send.shape[1:]), dtype=send.dtype, device=send.device)
        send = torch.cat([send, pad], 0)
    recv = torch.empty((world * biggest,) + tuple(send.shape[1:]), dtype=send.dtype, device=send.device)
    dist.all_gather_into_tensor(recv, send, group=group)
    if n_total == world * biggest:
        return recv
    parts = []
    for r in range(world):
        rb, re_ = shard_range(n_total, r, world)
        parts.append(recv[r * biggest: r * biggest + (re_ - rb)])
    return torch.cat(parts, 0)


def sample_sharded(sample_fn: Callable[..., torch.Tensor], y: torch.Tensor, vals: torch.Tensor,
                   mask: torch.Tensor, *, seed: int, group=None, gather: bool = True,
                   x_T: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Run ``sample_fn(y, vals, mask, seed=, sample_offset=, x_T=)`` on this rank's contiguous
    shard of the GLOBAL (y, vals, mask[, x_T]) and all-gather the outputs.

    ``sample_fn`` is normally a closure over ``Diffuser.sample_latent_cond(..., return_u8=True)``
    (see bench.py); it must return a tensor whose dim 0 is the shard size."""
    n = int(y.shape[0])
    if vals.shape[0] != n or mask.shape[0] != n:
        raise ValueError("y, vals and mask must describe the same global batch")
    if dist.is_available() and dist.is_initialized():
        world, rank = dist.get_world_size(group), dist.get_rank(group)
    else:
        world, rank = 1, 0
    b, e = shard_range(n, rank, world)
    out = sample_fn(y[b:e], vals[b:e], mask[b:e], seed=seed, sample_offset=b,
                    x_T=None if x_T is None else x_T[b:e])
    return gather_shards(out, n, group) if gather else out
