"""Vocab-parallel logprob / entropy / loss: lm_head rows sharded over the ranks of one box (BASELINE config 4).

Not in the reference (it always holds the whole ``lm_head``): a new capability the
north-star asks for.  Layout: rank r owns the contiguous vocab rows [r*ceil(V/R), ...); every
rank holds the same hidden [M,H] and targets.  Forward: K1 reduces the local rows to
(max, sum-exp, sum e*z, z_target|0) per token -- 16 B/token/rank -- ONE all-gather over
NCCL/NVLink moves them, and the finalize kernel merges the R partials in rank order (so every
rank gets bit-identical logprobs and therefore identical K3 coefficients, no further
exchange).  Backward: dW_r is local; dH = sum_r dlogits_r . W_r is one all-reduce.
"""
from __future__ import annotations

import torch
import torch.distributed as dist


def shard_range(vocab: int, world: int, rank: int, multiple: int = 8) -> tuple[int, int]:
    """Contiguous [begin, end) of vocab rows owned by ``rank``; shard size rounded up to ``multiple``."""
    per = -(-vocab // world)
    per = -(-per // multiple) * multiple
    b = min(rank * per, vocab)
    return b, min(b + per, vocab)


def exchange_stats(stats: torch.Tensor, group=None) -> torch.Tensor:
    """All-gather the per-rank [M,4] f32 partials into [R,M,4] (rank order), on the current stream."""
    world = dist.get_world_size(group) if (dist.is_available() and dist.is_initialized()) else 1
    if world == 1:
        return stats[None]
    flat = torch.empty(world * stats.shape[0], *stats.shape[1:], dtype=stats.dtype, device=stats.device)
    dist.all_gather_into_tensor(flat, stats.contiguous(), group=group)  # rank-major concatenation
    return flat.view(world, *stats.shape)


def reduce_grad_hidden(grad_hidden: torch.Tensor, group=None) -> torch.Tensor:
    """dH = sum over vocab shards (in place)."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(grad_hidden, group=group)
    return grad_hidden


def reduce_grad_hidden_async(grad_hidden: torch.Tensor, group=None):
    """Start dH = sum over vocab shards on NCCL's own stream (ordered after the work already queued on the current
    stream); returns a handle whose ``wait()`` orders the current stream after the reduction, or None for one rank."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        return dist.all_reduce(grad_hidden, group=group, async_op=True)
    return None


def shard_weight(weight: torch.Tensor, world: int, rank: int) -> tuple[torch.Tensor, int]:
    """(this rank's rows as a contiguous tensor, v_offset)."""
    b, e = shard_range(weight.shape[0], world, rank)
    return weight[b:e].contiguous(), b
