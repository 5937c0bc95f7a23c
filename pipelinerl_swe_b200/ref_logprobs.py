"""Reference-policy log-prob annotation with the forward-only fused lm_head.

Reference behaviour replaced: ``batch_annotate_traces_with_ref_logprobs`` (pipelinerl/preprocess.py:73-91)
posts every trace's prompt / completion token ids to a SEPARATE vLLM server holding the frozen reference
model (``llm.get_batch_logprobs_token_ids``) and copies the returned per-completion-token log-probs into
``trace["ref_logprobs"]``; those servers need their own GPUs (``preprocessor_fraction``,
pipelinerl/launch.py:47-48, world.py:143-145).  Here the same numbers come from one forward pass of a frozen
copy of the policy trunk on the trainer's GPUs: traces are packed into rows of at most ``max_tokens`` tokens
(position ids restart per trace, as in ``collate_packed``), the model runs under ``torch.no_grad()`` with its
``lm_head`` swapped for K1 -- 2*T*H*V FLOP, no backward, no [T, V] logits in HBM -- and the log-prob of
completion token ``t`` is read at position ``t - 1``.

Same contract as the reference: in-place update, ``len(trace["ref_logprobs"]) == len(trace["logprobs"])``,
sampling temperature 1 (vLLM reports log-probs of the unscaled distribution).
"""
from __future__ import annotations

from typing import Any, Sequence

import torch

from . import _abi
from .lm_head import LogprobRequest, install_fused_lm_head, logprob_request


def plan_packs(lengths: Sequence[int], max_tokens: int) -> list[list[int]]:
    """First-fit-in-order packing: consecutive traces share a row while the row stays <= max_tokens.

    A trace longer than ``max_tokens`` gets a row of its own (the model decides whether it can take it)."""
    packs: list[list[int]] = []
    cur: list[int] = []
    used = 0
    for i, n in enumerate(lengths):
        if n <= 0:
            raise ValueError(f"trace {i} has no tokens")
        if cur and used + n > max_tokens:
            packs.append(cur)
            cur, used = [], 0
        cur.append(i)
        used += n
    if cur:
        packs.append(cur)
    return packs


def completion_spans(traces: Sequence[dict[str, Any]]) -> list[tuple[int, int]]:
    """(n_tokens, n_completion) per trace; the completion is the LAST len(trace["logprobs"]) tokens (:77-79)."""
    out = []
    for i, tr in enumerate(traces):
        n, c = len(tr["input_ids"]), len(tr["logprobs"])
        if c >= n:
            raise ValueError(f"trace {i}: {c} completion tokens need a non-empty prompt ({n} tokens in total)")
        out.append((n, c))
    return out


@torch.no_grad()
def annotate_traces_with_ref_logprobs(model, traces: list[dict[str, Any]], max_tokens: int = 8192, temperature: float = 1.0,
                                      device: torch.device | str | None = None, group=None, v_offset: int = 0,
                                      pack: bool = True) -> list[dict[str, Any]]:
    """Fill ``trace["ref_logprobs"]`` for every trace from ``model`` (a frozen copy of the policy).  In place.

    ``pack=True`` relies, like the reference's packed training batches (finetune_loop.py:279-287), on an attention
    implementation that separates packed sequences by their restarting ``position_ids`` (flash-attention varlen);
    with any other attention pass ``pack=False``: one trace per forward row."""
    if not traces:
        return traces
    dev = torch.device(device) if device is not None else next(model.parameters()).device
    if dev.type != "cuda":
        raise _abi.B200GrpoError("annotate_traces_with_ref_logprobs: the model must live on a CUDA device (no CPU path)")
    install_fused_lm_head(model)
    spans = completion_spans(traces)
    was_training = model.training
    model.eval()
    try:
        for members in plan_packs([n for n, _ in spans], max_tokens if pack else 0):
            ids = torch.tensor([t for i in members for t in traces[i]["input_ids"]], dtype=torch.long).unsqueeze(0)
            pos = torch.cat([torch.arange(spans[i][0]) for i in members]).unsqueeze(0)
            ids_d, pos_d = ids.to(dev, non_blocking=True), pos.to(dev, non_blocking=True)
            with logprob_request(LogprobRequest(ids_d, float(temperature), group, v_offset)):
                out = model(input_ids=ids_d, attention_mask=torch.ones_like(ids_d), position_ids=pos_d)
            lp_ent = out.logits
            if lp_ent.shape[-1] != 2:
                raise _abi.B200GrpoError("model did not route through FusedLMHead (logits last dim != 2)")
            lp = lp_ent[0, :, 0].float().cpu()  # position p scores token p + 1
            off = 0
            for i in members:
                n, c = spans[i]
                traces[i]["ref_logprobs"] = lp[off + n - c - 1: off + n - 1].tolist()
                assert len(traces[i]["ref_logprobs"]) == len(traces[i]["logprobs"]), (
                    f"{len(traces[i]['ref_logprobs'])} != {len(traces[i]['logprobs'])}")  # :88-90
                off += n
    finally:
        model.train(was_training)
    return traces
