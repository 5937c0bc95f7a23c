"""Vocab-parallel logprob / entropy / loss: lm_head rows sharded over the ranks of one box (BASELINE config 4).

Not in the reference (it always holds the whole ``lm_head``): a new capability the
north-star asks for.  Layout: rank r owns the contiguous vocab rows [r*ceil(V/R), ...); every
rank holds the same hidden [M,H] and targets.  Forward: K1 reduces the local rows to
(max, sum-exp, sum e*z, z_target|0) per token -- 16 B/token/rank -- ONE all-gather over
NCCL/NVLink moves them, and the finalize kernel merges the R partials in rank order (so every
rank gets bit-identical logprobs and therefore identical K3 coefficients, no further
exchange).  Backward: dW_r is local; dH = sum_r dlogits_r . W_r is one all-reduce.

Two transports.  Passing a ``torch.distributed`` group uses NCCL for both steps.  Passing a ``SymmetricExchange`` keeps
NCCL out of the path: the forward all-gather is the K1 merge kernel's own ``multimem.st`` into a symmetric buffer, and the
dH all-reduce is ``b200grpo_mc_allreduce_bf16`` (``multimem.ld_reduce`` / ``multimem.st`` through the NVSwitch) run beside
the dW GEMM on a side stream -- its CTAs use no shared memory, so they share SMs with the GEMM instead of displacing it.
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


def shard_offset(n_local_rows: int, group=None) -> int:
    """First vocab row of this rank's shard = the summed row counts of the lower ranks (one small all-gather; called once,
    when the sharded head is installed).  ``group``: a process group or a ``SymmetricExchange``."""
    pg = getattr(group, "group", group)
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(pg) == 1:
        return 0
    sizes = [None] * dist.get_world_size(pg)
    dist.all_gather_object(sizes, int(n_local_rows), group=pg)
    return int(sum(sizes[: dist.get_rank(pg)]))


def shard_weight(weight: torch.Tensor, world: int, rank: int) -> tuple[torch.Tensor, int]:
    """(this rank's rows as a contiguous tensor, v_offset)."""
    b, e = shard_range(weight.shape[0], world, rank)
    return weight[b:e].contiguous(), b


class SymmetricExchange:
    """Vocab-parallel exchange over NVSwitch multicast (torch symmetric memory is only the allocator / rendezvous / barrier).

    Every rank of ``group`` constructs it collectively with the same ``max_rows`` / ``hidden``.  Buffers: two slots of
    [R][max_rows][4] f32 for the forward partials (alternating per call, so a fast rank's next forward cannot overwrite
    what a slow rank is still finalizing: it would first have to pass the next barrier) and one bf16 [max_rows, hidden]
    for dH.  Pass the object where a process group is accepted (``LMHeadLogprobFunction``, ``LogprobRequest.group``).
    """

    def __init__(self, group, device, max_rows: int, hidden: int, allreduce_ctas: int = 16):
        import torch.distributed._symmetric_memory as symm
        self.group = group if group is not None else dist.group.WORLD
        self.rank, self.world = dist.get_rank(self.group), dist.get_world_size(self.group)
        self.device = torch.device(device)
        self.max_rows, self.hidden, self.ctas = int(max_rows), int(hidden), int(allreduce_ctas)
        self.stats = symm.empty(2 * self.world * self.max_rows * 4, dtype=torch.float32, device=self.device)
        self.stats_h = symm.rendezvous(self.stats, self.group)
        self.dh = symm.empty(self.max_rows * self.hidden, dtype=torch.bfloat16, device=self.device)
        self.dh_h = symm.rendezvous(self.dh, self.group)
        for h in (self.stats_h, self.dh_h):
            if not h.has_multicast_support or not h.multicast_ptr:
                raise RuntimeError("SymmetricExchange: this group has no NVSwitch multicast mapping; pass the process group (NCCL) instead")
        self.parity = 0
        self.side = torch.cuda.Stream(self.device)

    # ---- forward: partials land in every rank's copy of the slot; returns the local [R, M, 4] view after the barrier
    def forward_stats(self, run_k1, M: int) -> torch.Tensor:
        """``run_k1(stats_mc)`` must launch K1 with that multicast slot address (``ops.lmhead_fwd_stats(..., stats_mc=...)``)."""
        assert M <= self.max_rows
        p, self.parity = self.parity, self.parity ^ 1
        base = p * self.world * self.max_rows * 4
        run_k1(int(self.stats_h.multicast_ptr) + 4 * (base + self.rank * M * 4))
        self.stats_h.barrier(channel=0)
        return self.stats[base: base + self.world * M * 4].view(self.world, M, 4)

    # ---- backward: the dH GEMM writes into dh_view(); allreduce_async() sums it over ranks on the side stream
    def dh_view(self, M: int) -> torch.Tensor:
        assert M <= self.max_rows
        return self.dh[: M * self.hidden].view(M, self.hidden)

    def allreduce_dh_async(self, M: int) -> torch.cuda.Event:
        main = torch.cuda.current_stream(self.device)
        self.side.wait_stream(main)                      # the dH GEMM on the main stream has been issued
        with torch.cuda.stream(self.side):
            self.dh_h.barrier(channel=1)                 # every rank's partial is complete
            from . import ops
            ops.mc_allreduce_bf16(int(self.dh_h.multicast_ptr), M * self.hidden, self.rank, self.world, self.ctas)
            self.dh_h.barrier(channel=1)                 # every slice has been written back everywhere
            done = torch.cuda.Event()
            done.record(self.side)
        return done
