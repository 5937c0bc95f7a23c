"""lm_head projection fused with log-softmax / entropy / gather (K1) and its backward (K2).

Replaces ``model(**inputs).logits`` -> ``/temperature`` -> ``log_softmax`` -> ``softmax`` ->
entropy -> ``gather`` of pipelinerl/finetune/rl/__init__.py:197-208: the [tokens x vocab]
logits never exist in HBM.  ``FusedLMHead`` is a drop-in for ``model.lm_head``: it owns the
same ``weight`` Parameter, so ZeRO-3 / FSDP gather hooks and tied embeddings keep working,
and only switches to the fused op while ``rl_step`` has an active ``LogprobRequest``.
"""
from __future__ import annotations

import threading
from dataclasses import dataclass

import torch
from torch import nn

from . import _abi, ops
from .vocab_parallel import SymmetricExchange, exchange_stats, reduce_grad_hidden, reduce_grad_hidden_async


_SAVE_LOGITS = True


def save_logits_enabled() -> bool:
    return _SAVE_LOGITS


def set_save_logits(flag: bool) -> None:
    """True (default): forward keeps bf16 logits (2 B x tokens x vocab until backward), backward skips the
    recompute GEMM.  False: nothing of size tokens x vocab is kept; backward recomputes the logits."""
    global _SAVE_LOGITS
    _SAVE_LOGITS = bool(flag)


_SKIP_MASKED = False


def skip_masked_tokens_enabled() -> bool:
    return _SKIP_MASKED


def set_skip_masked_tokens(flag: bool) -> None:
    """Opt-in: compute the lm_head path only for positions whose next token is labelled.

    The reference projects every position to the vocabulary and then multiplies everything it derived from
    them by ``labels != -100`` (rl/__init__.py:152-157, rl/utils.py:25-30): the log-prob, entropy and gradient
    of a position whose next token is a prompt / observation token never reach the loss or a statistic.  With
    this on, those rows are left out of the three GEMMs (their outputs read 0), so a batch with a fraction f of
    masked tokens costs (1 - f) of the FLOPs; labelled positions get bit-identical log-probs.  Costs one extra
    host sync per step (the number of labelled rows sizes the launch).  Off by default: exact drop-in."""
    global _SKIP_MASKED
    _SKIP_MASKED = bool(flag)


class LMHeadLogprobFunction(torch.autograd.Function):
    """(hidden [M,H] bf16, weight [Vloc,H] bf16, targets [M] i64) -> out f32 [M,2] = (logprob, entropy).

    With a process ``group`` the vocab rows are sharded contiguously over its ranks
    (vocab-parallel): each rank reduces its rows to (max, sumexp, sum e*z, z_target), the
    [M,4] stats are all-gathered over NCCL and merged; backward all-reduces grad_hidden.
    """

    @staticmethod
    def forward(ctx, hidden, weight, targets, inv_temperature: float, group=None, v_offset: int = 0):
        # keep bf16 logits for a recompute-free backward when a backward will follow
        need_bwd = ctx.needs_input_grad[0] or ctx.needs_input_grad[1]
        saved = ops.lmhead_saved_buffer(hidden.shape[0], hidden.shape[1], weight.shape[0], hidden.device) \
            if (need_bwd and save_logits_enabled()) else None
        if isinstance(group, SymmetricExchange):  # the merge kernel's multimem.st IS the all-gather (NVSwitch multicast)
            stats = group.forward_stats(lambda mc: ops.lmhead_fwd_stats(hidden, weight, targets, inv_temperature, v_offset, saved, stats_mc=mc),
                                        hidden.shape[0])
        else:
            stats = ops.lmhead_fwd_stats(hidden, weight, targets, inv_temperature, v_offset, saved)
            if group is not None:
                stats = exchange_stats(stats, group)  # [R,M,4], 16 B/token/rank over NCCL
        logprob, entropy, lse = ops.lmhead_finalize(stats)
        ctx.save_for_backward(hidden, weight, targets, lse, entropy)
        ctx.inv_temperature, ctx.group, ctx.v_offset = inv_temperature, group, v_offset
        ctx.saved = saved
        return torch.stack((logprob, entropy), dim=1)

    @staticmethod
    def backward(ctx, grad_out):
        hidden, weight, targets, lse, entropy = ctx.saved_tensors
        g = grad_out.float()
        coef_lp, coef_ent = g[:, 0].contiguous(), g[:, 1].contiguous()
        need_h, need_w = ctx.needs_input_grad[0], ctx.needs_input_grad[1]
        saved, ctx.saved = ctx.saved, None  # consumed in place by the backward; a second backward recomputes
        args = (hidden, weight, targets, lse, entropy, coef_lp, coef_ent, ctx.inv_temperature, ctx.v_offset)
        if isinstance(ctx.group, SymmetricExchange) and need_h:
            # dH partial straight into the symmetric buffer; in-switch all-reduce on a side stream WHILE this rank's dW GEMM runs
            x, M = ctx.group, hidden.shape[0]
            ops.lmhead_bwd(*args, need_grad_weight=False, saved=saved, stages=_abi.BWD_DLOGITS | _abi.BWD_DH, grad_hidden_out=x.dh_view(M))
            done = x.allreduce_dh_async(M)
            gw = ops.lmhead_bwd(*args, need_grad_hidden=False, saved=saved, stages=_abi.BWD_DW)[1] if need_w else None
            torch.cuda.current_stream().wait_event(done)
            return x.dh_view(M).clone(), gw, None, None, None, None   # the symmetric buffer is reused by the next micro-batch
        if ctx.group is not None and need_h and need_w:
            # vocab-parallel: dH = sum over shards is all-reduced on NCCL's stream WHILE this rank's dW GEMM runs
            gh, _ = ops.lmhead_bwd(*args, need_grad_weight=False, saved=saved, stages=_abi.BWD_DLOGITS | _abi.BWD_DH)
            pending = reduce_grad_hidden_async(gh, ctx.group)
            _, gw = ops.lmhead_bwd(*args, need_grad_hidden=False, saved=saved, stages=_abi.BWD_DW)
            if pending is not None:
                pending.wait()
            return gh, gw, None, None, None, None
        gh, gw = ops.lmhead_bwd(*args, need_grad_hidden=need_h, need_grad_weight=need_w, saved=saved)
        if gh is not None and ctx.group is not None:
            reduce_grad_hidden(gh, ctx.group)  # dH = sum over vocab shards
        return gh, gw, None, None, None, None


def lmhead_logprobs(hidden: torch.Tensor, weight: torch.Tensor, input_ids: torch.Tensor, temperature: float, group=None,
                    v_offset: int = 0, labels: torch.Tensor | None = None) -> torch.Tensor:
    """hidden [B,L,H], input_ids [B,L] -> f32 [B,L,2]: (logprob of input_ids[t+1], entropy) per position.

    Position L-1 of every row has no next token: it is scored against token 0 and the
    caller drops it (the reference slices ``[:, :-1]``, rl/__init__.py:200,208).
    With ``labels`` and ``set_skip_masked_tokens(True)`` only positions whose next token is labelled are computed.
    """
    B, L, H = hidden.shape
    targets = torch.roll(input_ids, shifts=-1, dims=1).reshape(-1).contiguous()
    h2 = hidden.reshape(B * L, H)
    if h2.dtype != torch.bfloat16:
        h2 = h2.to(torch.bfloat16)
    w = weight if weight.dtype == torch.bfloat16 else weight.to(torch.bfloat16)
    inv_t = 1.0 / float(temperature)
    if labels is not None and skip_masked_tokens_enabled():
        keep = torch.roll(labels, shifts=-1, dims=1) != -100
        keep[:, -1] = False                                  # no next token
        idx = keep.reshape(-1).nonzero().squeeze(1)          # host sync: the row count sizes the launch
        out = torch.zeros(B * L, 2, dtype=torch.float32, device=hidden.device)
        if idx.numel() > 0:
            part = LMHeadLogprobFunction.apply(h2.index_select(0, idx), w.contiguous(), targets.index_select(0, idx), inv_t, group,
                                               v_offset)
            out = out.index_copy(0, idx, part)               # autograd scatters the gradient rows back (zeros elsewhere)
        return out.reshape(B, L, 2)
    out = LMHeadLogprobFunction.apply(h2.contiguous(), w.contiguous(), targets, inv_t, group, v_offset)
    return out.reshape(B, L, 2)


@dataclass
class LogprobRequest:
    """What ``rl_step`` asks the swapped lm_head to compute during ``model(**inputs)``."""
    input_ids: torch.Tensor
    temperature: float
    group: object = None
    v_offset: int = 0
    labels: torch.Tensor | None = None   # only read when set_skip_masked_tokens(True)


_tls = threading.local()


class logprob_request:
    """Context manager activating a LogprobRequest for FusedLMHead on this thread."""

    def __init__(self, req: LogprobRequest):
        self.req = req

    def __enter__(self):
        _tls.req = self.req
        return self.req

    def __exit__(self, *exc):
        _tls.req = None
        return False


class FusedLMHead(nn.Module):
    """``model.lm_head`` (an ``nn.Linear(H, V, bias=False)``) with the fused forward.

    Outside ``rl_step`` it is the plain projection (generation / evaluation unchanged).
    Inside, ``forward(hidden)`` returns [B, L, 2] (logprob, entropy) in place of logits.

    ``install_fused_lm_head`` does not build a new module: it re-classes the existing one IN PLACE to a subclass
    of (FusedLMHead, its own class), so the module object, its ``weight`` Parameter, its name in the state dict and
    every hook already registered on it (ZeRO-3 parameter gathering, FSDP, DDP, profilers) stay exactly as they were.
    ``FusedLMHead(linear)`` builds a stand-alone wrapper around the same Parameter for code that wants one.
    """

    def __init__(self, linear: nn.Module):
        super().__init__()
        _check_fusable(linear)
        self.weight = linear.weight  # same Parameter object: tied embeddings stay tied
        self.in_features = linear.weight.shape[1]
        self.out_features = linear.weight.shape[0]

    def forward(self, hidden: torch.Tensor) -> torch.Tensor:
        req = getattr(_tls, "req", None)
        if req is None:
            if type(self) is FusedLMHead:
                return nn.functional.linear(hidden, self.weight)
            return super().forward(hidden)  # the original class's own forward (plain Linear, LoRA-wrapped, ...)
        return lmhead_logprobs(hidden, self.weight, req.input_ids, req.temperature, req.group, req.v_offset, req.labels)


def _check_fusable(linear: nn.Module) -> None:
    if getattr(linear, "bias", None) is not None:
        raise ValueError("FusedLMHead supports bias-free lm_head only (all Qwen2.5 / Llama heads)")
    if not isinstance(getattr(linear, "weight", None), torch.Tensor) or linear.weight.dim() != 2:
        raise ValueError("lm_head must own a 2-D `weight`")


def install_fused_lm_head(model: nn.Module) -> FusedLMHead:
    """Idempotently give ``model.lm_head`` the fused forward, looking through the usual wrappers: DeepSpeed / DDP
    ``.module``, the reference's value-head wrapper ``.pretrained_model`` (finetune/value_model.py:60-66), PEFT
    ``.base_model`` / ``.model``."""
    owner = model
    for _ in range(6):
        if hasattr(owner, "lm_head"):
            break
        nxt = None
        for attr in ("module", "pretrained_model", "base_model", "model"):
            cand = getattr(owner, attr, None)
            if isinstance(cand, nn.Module) and cand is not owner:
                nxt = cand
                break
        if nxt is None:
            break
        owner = nxt
    head = getattr(owner, "lm_head", None)
    if head is None:
        raise AttributeError("rl_step needs a model with an `lm_head` projection to fuse")
    if not isinstance(head, FusedLMHead):
        _check_fusable(head)
        base = type(head)
        head.__class__ = type("Fused" + base.__name__, (FusedLMHead, base), {"__module__": __name__})
    return head
