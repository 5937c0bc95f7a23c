"""lm_head projection fused with log-softmax / entropy / gather (K1) and its backward (K2).

Replaces ``model(**inputs).logits`` -> ``/temperature`` -> ``log_softmax`` -> ``softmax`` ->
entropy -> ``gather`` of pipelinerl/finetune/rl/__init__.py:197-208.  No fp32 logits / log-probs / probs tensor is
ever formed.  What IS kept between forward and backward depends on the mode: by default (``set_save_logits(True)``) the
forward stores bf16 logits relative to their tile maximum -- 2 B x tokens x vocab, 2.49 GB at T=8192, capped by
``set_saved_logits_budget`` -- so that the backward needs no third GEMM; with ``set_save_logits(False)`` nothing of
size tokens x vocab survives the forward and the backward recomputes the logits tile by tile (the north-star design,
~20 % slower).  Either way the backward's bf16 dlogits scratch is bounded by the same budget (row chunks).
``FusedLMHead`` is a drop-in for ``model.lm_head``: it owns the
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
_SAVED_BUDGET = 4 << 30
_FUSED_GRAD_ACC = False


def save_logits_enabled() -> bool:
    return _SAVE_LOGITS


def set_save_logits(flag: bool) -> None:
    """True (default): forward keeps bf16 logits (2 B x tokens x vocab until backward, capped by
    ``set_saved_logits_budget``), backward skips the recompute GEMM.  False: nothing of size tokens x vocab is kept
    between forward and backward; backward recomputes the logits tile by tile."""
    global _SAVE_LOGITS
    _SAVE_LOGITS = bool(flag)


def saved_logits_budget() -> int:
    return _SAVED_BUDGET


def set_saved_logits_budget(nbytes: int) -> None:
    """Upper bound (bytes, default 4 GiB) on every [rows x vocab]-sized bf16 buffer of the lm_head path: the logits the
    forward keeps AND the dlogits scratch of the backward.  A micro-batch whose rows x vocab x 2 B exceeds it is cut into
    row chunks that fit (whole multiples of 512 rows, at least 4096 so the dW accumulate stays amortised): the first chunk
    keeps its logits, the others are recomputed in the backward, one chunk at a time through the same scratch.  A 50 000
    token row (the reference's SWE recipe, launch_train.sh:3) then needs 4 GiB instead of 15.2 GB, at recompute speed for
    the rows beyond the first chunk."""
    global _SAVED_BUDGET
    _SAVED_BUDGET = int(nbytes)


def fused_grad_accumulation_enabled() -> bool:
    return _FUSED_GRAD_ACC


def set_fused_grad_accumulation(flag: bool) -> None:
    """Opt-in (Megatron ``main_grad`` convention).  Off: backward returns a fresh bf16 [V,H] dW and autograd adds it to
    ``lm_head.weight.grad`` in bf16, as for any module.  On: the dW GEMM's epilogue adds straight into the f32 buffer
    ``lm_head.weight.main_grad`` (``red.global.add.v4.f32``; allocated zeroed on first use) and autograd sees no gradient
    for the weight -- no 2 B x V x H temporary per micro-batch, f32 accumulation across the micro-batches of an optimizer
    step.  The optimizer side reads ``weight.main_grad`` or calls ``fold_main_grad(model)`` before its step.

    Precondition: nothing in the trainer may wait for an autograd gradient on ``lm_head.weight`` -- torch DDP's reducer, DeepSpeed
    ZeRO's reduce hooks and FSDP do; with those keep this off (the default) or fold before THEIR reduction point."""
    global _FUSED_GRAD_ACC
    _FUSED_GRAD_ACC = bool(flag)


def fold_main_grad(model: nn.Module, zero: bool = True) -> None:
    """Add every ``main_grad`` written by the fused accumulate into the parameter's ``.grad`` (created if absent) in the
    parameter's dtype, and zero the f32 buffer: what an unmodified optimizer / grad-clipping loop expects to find."""
    for p in model.parameters():
        mg = getattr(p, "main_grad", None)
        if mg is None:
            continue
        if p.grad is None:
            p.grad = mg.to(p.dtype, copy=True)   # copy: an f32 parameter would otherwise alias the buffer zeroed below
        else:
            p.grad.add_(mg.to(p.dtype))
        if zero:
            mg.zero_()


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


def row_chunks(M: int, Vloc: int, budget: int | None = None) -> list[tuple[int, int]]:
    """Row ranges of one micro-batch such that rows x ceil(Vloc/256)*256 x 2 B <= budget per chunk (see
    ``set_saved_logits_budget``).  One chunk when everything fits."""
    budget = saved_logits_budget() if budget is None else budget
    per_row = -(-Vloc // 256) * 256 * 2
    fit = max(4096, budget // per_row // 512 * 512)
    if M <= fit:
        return [(0, M)]
    n = -(-M // fit)
    size = min(fit, -(-(-(-M // n)) // 512) * 512)     # balanced: no sliver of a last chunk
    return [(r, min(M, r + size)) for r in range(0, M, size)]


class _GradSink:
    """Carries the weight Parameter into the autograd Function without making it a differentiable input twice."""

    def __init__(self, param, force_w: bool = False):
        self.param = param
        self.force_w = force_w    # the operand handed to the Function is a detached cast of ``param``: dW must go to main_grad


class LMHeadLogprobFunction(torch.autograd.Function):
    """(hidden [M,H] bf16, weight [Vloc,H] bf16, targets [M] i64) -> out f32 [M,2] = (logprob, entropy).

    With a process ``group`` the vocab rows are sharded contiguously over its ranks
    (vocab-parallel): each rank reduces its rows to (max, sumexp, sum e*z, z_target), the
    [M,4] stats are all-gathered over NCCL and merged; backward all-reduces grad_hidden.
    """

    @staticmethod
    def forward(ctx, hidden, weight, targets, inv_temperature: float, group=None, v_offset: int = 0, sink: _GradSink | None = None,
                entropy_grad: bool = True):
        M, H = hidden.shape
        Vloc = weight.shape[0]
        # an fp32 (or fp16) master weight under fused accumulation: the bf16 operand is a detached cast, the f32 gradient goes
        # straight to the master's main_grad through the sink, autograd sees no weight gradient at all
        ctx.sink_forces_w = bool(sink is not None and sink.force_w)
        need_bwd = ctx.needs_input_grad[0] or ctx.needs_input_grad[1] or ctx.sink_forces_w
        chunks = row_chunks(M, Vloc)
        # keep bf16 logits for a recompute-free backward when a backward will follow: the whole micro-batch when it fits
        # the budget, else its first row chunk (the other chunks are recomputed in the backward)
        # An entropy gradient (entropy_bonus != 0) is always formed from recomputed f32 logits: on near-deterministic rows
        # p * (log p + H) at the peak token is a difference of nearly equal numbers, and the 2^-9 of a bf16 exponential ~ 1 is as
        # large as H itself (found by tests/fuzz_rl_step.py: 30 % gradient error on a row of entropy 4.5e-3 nats).
        saved = None
        if need_bwd and save_logits_enabled() and not entropy_grad:
            r0, r1 = chunks[0]
            saved = ops.lmhead_saved_buffer(r1 - r0, H, Vloc, hidden.device)

        def k1(stats_mc=None):
            """K1 over all rows -> [M,4] partials (or stored through the multicast address)."""
            if len(chunks) == 1:
                return ops.lmhead_fwd_stats(hidden, weight, targets, inv_temperature, v_offset, saved, stats_mc=stats_mc)
            parts = []
            for ci, (r0, r1) in enumerate(chunks):
                mc = None if stats_mc is None else stats_mc + 16 * r0
                parts.append(ops.lmhead_fwd_stats(hidden[r0:r1], weight, targets[r0:r1], inv_temperature, v_offset,
                                                  saved if ci == 0 else None, stats_mc=mc))
            return None if stats_mc is not None else torch.cat(parts)

        if isinstance(group, SymmetricExchange):  # the merge kernel's multimem.st IS the all-gather (NVSwitch multicast)
            stats = group.forward_stats(k1, M)
        else:
            stats = k1()
            if group is not None:
                stats = exchange_stats(stats, group)  # [R,M,4], 16 B/token/rank over NCCL
        logprob, entropy, lse = ops.lmhead_finalize(stats)
        ctx.save_for_backward(hidden, weight, targets, lse, entropy)
        ctx.inv_temperature, ctx.group, ctx.v_offset = inv_temperature, group, v_offset
        ctx.saved, ctx.chunks, ctx.sink, ctx.entropy_grad = saved, chunks, sink, bool(entropy_grad)
        return torch.stack((logprob, entropy), dim=1)

    @staticmethod
    def backward(ctx, grad_out):
        hidden, weight, targets, lse, entropy = ctx.saved_tensors
        g = grad_out.float()
        # entropy_grad False (rl_step with entropy_bonus == 0: K3's dLoss/dentropy is identically zero): the backward runs
        # its GEMMs straight on the saved exponentials, no [tokens x vocab] dlogits pass
        coef_lp, coef_ent = g[:, 0].contiguous(), (g[:, 1].contiguous() if ctx.entropy_grad else None)
        need_h, need_w = ctx.needs_input_grad[0], ctx.needs_input_grad[1] or ctx.sink_forces_w
        saved, ctx.saved = ctx.saved, None  # consumed in place by the backward; a second backward recomputes
        chunks, group = ctx.chunks, ctx.group
        M, H = hidden.shape
        symmetric = isinstance(group, SymmetricExchange)
        # ---- where the gradients go
        gh = None
        if need_h:   # vocab-parallel over multicast: the dH partial is written straight into the symmetric buffer
            gh = group.dh_view(M) if symmetric else torch.empty(M, H, dtype=torch.bfloat16, device=hidden.device)
        gw, gw_ret, acc_first = None, None, False
        if need_w:
            sink = ctx.sink
            if sink is not None and (fused_grad_accumulation_enabled() or ctx.sink_forces_w):
                p = sink.param
                if getattr(p, "main_grad", None) is None:
                    p.main_grad = torch.zeros(weight.shape, dtype=torch.float32, device=weight.device)
                gw, acc_first = p.main_grad, True           # f32 accumulate in the GEMM epilogue; autograd gets no dW
            elif len(chunks) > 1:
                gw = torch.empty(weight.shape, dtype=torch.float32, device=weight.device)   # chunks accumulate in f32
            else:
                gw = torch.empty(weight.shape, dtype=torch.bfloat16, device=weight.device)
        # ---- row chunks (one unless the micro-batch exceeds the [rows x vocab] budget); the collective on dH is started
        #      after the LAST chunk's dH so that it runs beside that chunk's dW GEMM
        overlap = group is not None and need_h and need_w
        pending = done = None
        for ci, (r0, r1) in enumerate(chunks):
            a = (hidden[r0:r1], weight, targets[r0:r1], lse[r0:r1], entropy[r0:r1], coef_lp[r0:r1],
                 None if coef_ent is None else coef_ent[r0:r1], ctx.inv_temperature, ctx.v_offset)
            kw = dict(saved=saved if ci == 0 else None, grad_hidden_out=gh[r0:r1] if need_h else None, grad_weight_out=gw)
            acc = need_w and (acc_first or ci > 0)
            if overlap and ci == len(chunks) - 1:
                ops.lmhead_bwd(*a, need_grad_weight=False, stages=_abi.BWD_DLOGITS | _abi.BWD_DH, **kw)
                if symmetric:   # in-switch all-reduce on a side stream WHILE this rank's dW GEMM runs
                    done = group.allreduce_dh_async(M)
                else:           # dH = sum over shards on NCCL's stream WHILE this rank's dW GEMM runs
                    pending = reduce_grad_hidden_async(gh, group)
                ops.lmhead_bwd(*a, need_grad_hidden=False, stages=_abi.BWD_DW, accumulate=acc, **kw)
            else:
                ops.lmhead_bwd(*a, need_grad_hidden=need_h, need_grad_weight=need_w, accumulate=acc, **kw)
            saved = kw = None   # the first chunk's logits are consumed: their memory serves the next chunk's dlogits scratch
        if done is not None:
            torch.cuda.current_stream().wait_event(done)
        elif pending is not None:
            pending.wait()
        elif group is not None and need_h:
            if symmetric:
                torch.cuda.current_stream().wait_event(group.allreduce_dh_async(M))
            else:
                reduce_grad_hidden(gh, group)  # dH = sum over vocab shards
        if symmetric and need_h:
            gh = gh.clone()   # the symmetric buffer is reused by the next micro-batch
        if need_w and not acc_first:
            gw_ret = gw if gw.dtype == torch.bfloat16 else gw.to(torch.bfloat16)
        return gh, gw_ret, None, None, None, None, None, None


def lmhead_logprobs(hidden: torch.Tensor, weight: torch.Tensor, input_ids: torch.Tensor, temperature: float, group=None,
                    v_offset: int = 0, labels: torch.Tensor | None = None, boundary_ids: torch.Tensor | None = None,
                    boundary_labels: torch.Tensor | None = None, entropy_grad: bool = True) -> torch.Tensor:
    """hidden [B,L,H], input_ids [B,L] -> f32 [B,L,2]: (logprob of input_ids[t+1], entropy) per position.

    Position L-1 of every row has no next token: it is scored against token 0 and the
    caller drops it (the reference slices ``[:, :-1]``, rl/__init__.py:200,208) -- unless ``boundary_ids`` [B] names
    the token that follows each row (the first token of the next sequence-parallel slice): then L-1 is scored against it.
    With ``labels`` and ``set_skip_masked_tokens(True)`` only positions whose next token is labelled are computed.
    """
    B, L, H = hidden.shape
    targets = torch.roll(input_ids, shifts=-1, dims=1)
    if boundary_ids is not None:
        targets[:, -1] = boundary_ids.to(targets.dtype)
    targets = targets.reshape(-1).contiguous()
    h2 = hidden.reshape(B * L, H)
    if h2.dtype != torch.bfloat16:
        h2 = h2.to(torch.bfloat16)
    inv_t = 1.0 / float(temperature)
    is_param = isinstance(weight, nn.Parameter)
    if weight.dtype == torch.bfloat16:
        w = weight
        sink = _GradSink(weight) if is_param else None       # fused f32 accumulate target
    elif is_param and weight.requires_grad and hidden.requires_grad and fused_grad_accumulation_enabled() and torch.is_grad_enabled():
        # fp32 master weights (torch AMP / accelerate mixed precision): the GEMMs read a bf16 cast made per call -- what autocast
        # does for F.linear -- and dW is accumulated in f32 straight into weight.main_grad; no bf16 [V,H] gradient, no cast back
        w = weight.detach().to(torch.bfloat16)
        sink = _GradSink(weight, force_w=True)
    else:
        w = weight.to(torch.bfloat16)                        # autograd casts the bf16 gradient back to the parameter's dtype
        sink = None
    if labels is not None and skip_masked_tokens_enabled():
        keep = torch.roll(labels, shifts=-1, dims=1) != -100
        keep[:, -1] = False if boundary_labels is None else (boundary_labels != -100)   # no next token unless it is carried
        idx = keep.reshape(-1).nonzero().squeeze(1)          # host sync: the row count sizes the launch
        out = torch.zeros(B * L, 2, dtype=torch.float32, device=hidden.device)
        if idx.numel() > 0:
            part = LMHeadLogprobFunction.apply(h2.index_select(0, idx), w.contiguous(), targets.index_select(0, idx), inv_t, group,
                                               v_offset, sink, entropy_grad)
            out = out.index_copy(0, idx, part)               # autograd scatters the gradient rows back (zeros elsewhere)
        return out.reshape(B, L, 2)
    out = LMHeadLogprobFunction.apply(h2.contiguous(), w.contiguous(), targets, inv_t, group, v_offset, sink, entropy_grad)
    return out.reshape(B, L, 2)


@dataclass
class LogprobRequest:
    """What ``rl_step`` asks the swapped lm_head to compute during ``model(**inputs)``."""
    input_ids: torch.Tensor
    temperature: float
    group: object = None
    v_offset: int = 0
    labels: torch.Tensor | None = None   # only read when set_skip_masked_tokens(True)
    boundary_ids: torch.Tensor | None = None      # [B] next-slice first token (types.make_slices(carry_boundary_target=True))
    boundary_labels: torch.Tensor | None = None
    entropy_grad: bool = True                     # False: the caller guarantees dLoss/dentropy == 0 (entropy_bonus == 0)


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
        group, v_offset = req.group, req.v_offset
        if group is None and getattr(self, "vocab_group", None) is not None:   # sharded head declared at install time
            group, v_offset = self.vocab_group, self.v_offset
        return lmhead_logprobs(hidden, self.weight, req.input_ids, req.temperature, group, v_offset, req.labels,
                               req.boundary_ids, req.boundary_labels, req.entropy_grad)


def _check_fusable(linear: nn.Module) -> None:
    if getattr(linear, "bias", None) is not None:
        raise ValueError("FusedLMHead supports bias-free lm_head only (all Qwen2.5 / Llama heads)")
    if not isinstance(getattr(linear, "weight", None), torch.Tensor) or linear.weight.dim() != 2:
        raise ValueError("lm_head must own a 2-D `weight`")


def install_fused_lm_head(model: nn.Module, vocab_group=None, v_offset: int | None = None) -> FusedLMHead:
    """Idempotently give ``model.lm_head`` the fused forward, looking through the usual wrappers: DeepSpeed / DDP
    ``.module``, the reference's value-head wrapper ``.pretrained_model`` (finetune/value_model.py:60-66), PEFT
    ``.base_model`` / ``.model``.

    ``vocab_group`` (a process group, or a ``SymmetricExchange`` for the NVSwitch-multicast transport) declares that
    ``lm_head.weight`` holds only this rank's contiguous shard of the vocabulary rows (BASELINE config 4): every
    ``rl_step`` on this model then runs vocab-parallel.  ``v_offset`` = first vocab row of the shard; by default the summed
    shard sizes of the lower ranks (one all-gather at install time)."""
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
    if vocab_group is not None and (getattr(head, "vocab_group", None) is not vocab_group or v_offset is not None):
        from .vocab_parallel import shard_offset
        head.vocab_group = vocab_group
        head.v_offset = int(v_offset) if v_offset is not None else shard_offset(head.weight.shape[0], vocab_group)
    return head
