"""rl_step: drop-in for pipelinerl/finetune/rl/__init__.py:130-377 on hand-written sm_100a kernels.

Same signature and returns: ``rl_step(model, batch, current_step, max_step, config) ->
(loss, stats)`` with a 0-dim autograd-connected ``loss`` and the reference's 34 (39 with a
value head) float stats, or ``{"input_size": ...}`` when the batch holds no labels.
What differs is only how it is computed: K1 inside the swapped ``lm_head`` (no fp32
logits / log-probs / probs tensors; see lm_head.py for what the two backward modes keep), one K3 launch for every per-token term / reduction / metric, ONE host sync (the
metrics vector) instead of ~40, and K2 in backward.
"""
from __future__ import annotations

from typing import Any

import torch

from . import _abi
from ._abi import LossCfg
from . import ops
from .config import linear_decay_coef
from .lm_head import LogprobRequest, install_fused_lm_head, logprob_request

_POLICY = {"ppo": 0, "reinforce": 1}


class TokenLossFunction(torch.autograd.Function):
    """(logprob, entropy[, values]) -> loss; K3 also emits dLoss/d(inputs), so backward is a scale."""

    @staticmethod
    def forward(ctx, logprob, entropy, values, batch, cfg, sink):
        loss, metrics, coef_lp, coef_ent, coef_v = ops.token_loss(logprob.detach(), entropy.detach(), batch, cfg, values)
        sink["metrics"] = metrics
        ctx.save_for_backward(coef_lp, coef_ent, coef_v if coef_v is not None else coef_lp)
        ctx.has_values = coef_v is not None
        ctx.shape = logprob.shape
        return loss.reshape(())

    @staticmethod
    def backward(ctx, grad_loss):
        coef_lp, coef_ent, coef_v = ctx.saved_tensors
        g = grad_loss.float()
        gv = g * coef_v if ctx.has_values and ctx.needs_input_grad[2] else None
        return (g * coef_lp).reshape(ctx.shape), (g * coef_ent).reshape(ctx.shape), gv, None, None, None


def _batch_dict(batch: Any) -> dict:
    names = ("input_ids", "attention_mask", "labels", "position_ids", "rewards", "advantages", "ref_logprobs", "old_logprobs",
             "group_tokens", "num_labels", "overflow", "pixel_values", "image_grid_thw")
    get = (lambda k: batch.get(k)) if isinstance(batch, dict) else (lambda k: getattr(batch, k, None))
    d = {k: get(k) for k in names}
    d["is_packed"] = bool(get("is_packed"))
    d["boundary"] = get("boundary")   # opt-in SP-slice boundary target (types.make_slices(carry_boundary_target=True))
    return d


def make_loss_cfg(config: Any, current_step: int, max_step: int, is_packed: bool, has_values: bool) -> LossCfg:
    if config.policy_loss not in _POLICY:
        raise ValueError(f"Unknown algorithm {config.policy_loss}")  # rl/__init__.py:283
    return LossCfg(
        policy_loss=_POLICY[config.policy_loss], use_advantages=int(config.use_advantages),
        relu_log_p_weights=int(config.relu_log_p_weights), group_normalization=int(config.group_normalization),
        overlong_filtering=int(config.overlong_filtering), is_packed=int(is_packed), has_values=int(has_values),
        batch_size=int(config.batch_size), epsilon=float(config.epsilon),
        clamp_log_ratio_ref_new=float(config.clamp_log_ratio_ref_new_value),
        kl_coef=float(linear_decay_coef(current_step, max_step, config.kl_coef, config.final_kl_coef)),
        entropy_coef=float(linear_decay_coef(current_step, max_step, config.entropy_bonus, config.final_entropy_bonus)),
        value_loss_coef=float(config.value_loss_coef), reserved=0.0)


def stats_from_metrics(m: list[float], cfg: LossCfg, input_size: int, has_values: bool,
                       coefs: tuple[float, float] | None = None) -> dict[str, float]:
    """The reference's stats dict (rl/__init__.py:325-375), same keys in the same order."""
    A = _abi
    flags = int(m[A.M_FLAGS])
    # the reference's asserts, checked at the single D2H sync (:209,:222,:237,:264,:313)
    assert not flags & A.F_GROUP_TOKENS_LE0, "group_tokens must be greater than zero for group normalization"
    assert not flags & A.F_NEW_LP_NONFINITE, "new_logprobs is not finite"
    assert not flags & A.F_REF_NEW_NONFINITE, "log_ratio_ref_new is not finite"
    assert not flags & A.F_KL_NONFINITE, "approx_kl is not finite"
    assert not flags & A.F_LOSS_NONFINITE, f"Non-finite loss detected: {m[A.M_LOSS]}"
    if int(m[A.M_N_MASKED]) == 0:
        return {"input_size": float(input_size)}
    n_seq = int(m[A.M_N_SEQ])
    kl_coef, ent_coef = coefs if coefs is not None else (cfg.kl_coef, cfg.entropy_coef)
    loss = m[A.M_LOSS]
    stats = {
        "loss": loss, "max_loss": loss, "min_loss": loss,
        "reward": m[A.M_REWARD], "max_reward": m[A.M_MAX_REWARD], "min_reward": m[A.M_MIN_REWARD],
        "entropy": m[A.M_ENTROPY], "old_logprobs": m[A.M_OLD_LP], "new_logprobs": m[A.M_NEW_LP],
        "ref_logprobs": m[A.M_REF_LP],
        "advantage": m[A.M_ADVANTAGE], "max_advantage": m[A.M_MAX_ADV], "min_advantage": m[A.M_MIN_ADV],
        "kl": m[A.M_KL], "max_kl": m[A.M_MAX_KL], "min_kl": m[A.M_MIN_KL],
        "policy_loss": m[A.M_POLICY], "surr1": m[A.M_SURR1], "surr2": m[A.M_SURR2],
        "ratio_new_old": m[A.M_RATIO], "ratio_new_old_sum": m[A.M_RATIO_SUM],
        "ratio_new_old_squared_sum": m[A.M_RATIO_SQ_SUM],
        "ratio_ref_new": m[A.M_RATIO_REF_NEW], "ratio_ref_old": m[A.M_RATIO_REF_OLD],
        "clamp_log_ratio_ref_new_indicator": m[A.M_CLAMP_REF_NEW],
        "clamp_log_ratio_new_old_indicator": m[A.M_CLAMP_NEW_OLD],
        "num_nans": int(m[A.M_NUM_NANS]),
        "token_weight": m[A.M_TOKEN_WEIGHT], "max_token_weight": m[A.M_MAX_TW], "min_token_weight": m[A.M_MIN_TW],
        "kl_coef": n_seq * kl_coef, "entropy_bonus_coef": n_seq * ent_coef,
        "num_output_tokens_sum": int(m[A.M_N_MASKED]),
        "input_size": input_size,
    }
    if has_values:
        stats.update(value_mean=m[A.M_VALUE_MEAN], value_max=m[A.M_MAX_VALUE], value_min=m[A.M_MIN_VALUE],
                     value_loss=m[A.M_VALUE_LOSS], value_mse=m[A.M_VALUE_MSE])
    return stats


_PER_TOKEN = ("labels", "position_ids", "rewards", "advantages", "ref_logprobs", "old_logprobs", "group_tokens", "num_labels", "overflow")


def _append_boundary(b: dict) -> dict:
    """[1, L] fields + the next slice's first token -> [1, L+1], so that K3's shifted view ``[:, 1:]`` reaches the boundary
    target.  The carried position id is kept non-zero: the sequence that may start there is counted by the next slice."""
    bd = b["boundary"]
    out = dict(b)
    for k in _PER_TOKEN:
        if b.get(k) is None:
            continue
        extra = bd[k].to(b[k].dtype)
        if k == "position_ids":
            extra = extra.clamp(min=1)
        out[k] = torch.cat([b[k], extra], dim=1)
    return out


def grpo_loss_from_logprobs(lp_ent: torch.Tensor, b: dict, cfg: LossCfg, values: torch.Tensor | None = None):
    """lp_ent f32 [B,L,2] from the fused lm_head -> (loss, device metrics f64[40])."""
    if b.get("boundary") is not None:   # the last position was scored against the next slice's first token: keep it
        b = _append_boundary(b)
        if values is not None:
            values = torch.cat([values, values[:, -1:]], dim=1)   # placeholder for the carried position (never read as a prediction)
        logprob, entropy = lp_ent[:, :, 0].contiguous(), lp_ent[:, :, 1].contiguous()
    else:
        logprob = lp_ent[:, :-1, 0].contiguous()  # the reference's [:, :-1] (:200)
        entropy = lp_ent[:, :-1, 1].contiguous()
    sink: dict = {}
    loss = TokenLossFunction.apply(logprob, entropy, values, b, cfg, sink)
    return loss, sink["metrics"]


def _forward_loss(model, batch, current_step: int, max_step: int, config, vocab_group=None, v_offset=None):
    """Shared by rl_step / rl_step_async: model forward through the fused lm_head + K3.  No host sync."""
    b = _batch_dict(batch)
    has_value_head = hasattr(model, "value_head")
    cfg = make_loss_cfg(config, current_step, max_step, b["is_packed"], has_value_head)
    if not b["input_ids"].is_cuda:
        raise _abi.B200GrpoError("rl_step: batch must be on a CUDA device (no CPU path); call batch.to_device(...) first")
    head = install_fused_lm_head(model, vocab_group, v_offset)
    model_inputs = {"input_ids": b["input_ids"], "attention_mask": b["attention_mask"]}  # labels dropped: no HF CE pass
    if b["is_packed"]:
        model_inputs["position_ids"] = b["position_ids"]
    if b["pixel_values"] is not None:
        model_inputs["pixel_values"] = b["pixel_values"]
    if b["image_grid_thw"] is not None:
        model_inputs["image_grid_thw"] = b["image_grid_thw"]
    req = LogprobRequest(b["input_ids"], float(config.temperature), labels=b["labels"],
                         group=getattr(head, "vocab_group", None), v_offset=getattr(head, "v_offset", 0),
                         boundary_ids=None if b["boundary"] is None else b["boundary"]["input_ids"].reshape(-1),
                         boundary_labels=None if b["boundary"] is None else b["boundary"]["labels"].reshape(-1),
                         entropy_grad=cfg.entropy_coef != 0.0)   # K3's dLoss/dentropy = -w * entropy_coef: identically zero without a bonus
    with logprob_request(req):
        outputs = model(**model_inputs)
    lp_ent = outputs.logits
    if lp_ent.shape[-1] != 2:
        raise _abi.B200GrpoError("model did not route through FusedLMHead (logits last dim != 2)")
    values = outputs.value.float().contiguous() if has_value_head else None
    tb = {k: (v.contiguous() if isinstance(v, torch.Tensor) else v) for k, v in b.items()}
    loss, metrics = grpo_loss_from_logprobs(lp_ent.float(), tb, cfg, values)
    coefs = (linear_decay_coef(current_step, max_step, config.kl_coef, config.final_kl_coef),
             linear_decay_coef(current_step, max_step, config.entropy_bonus, config.final_entropy_bonus))
    return loss, metrics, cfg, coefs, b["input_ids"].numel(), has_value_head


def rl_step(model, batch, current_step: int, max_step: int, config, *, vocab_group=None,
            v_offset: int | None = None) -> tuple[torch.Tensor, dict[str, float]]:
    """Perform a single RL step; see module docstring.  Handles packed and unpacked batches.

    The five positional arguments are the reference's.  ``vocab_group`` (keyword-only, beyond the reference: BASELINE
    config 4) says ``model.lm_head.weight`` is this rank's contiguous shard of the vocabulary rows and names the ranks
    that hold the others -- a process group (NCCL exchange) or a ``SymmetricExchange`` (NVSwitch multicast kernels).
    Every rank must pass the same batch; each gets bit-identical log-probs, loss and stats.  Equivalent to calling
    ``install_fused_lm_head(model, vocab_group=...)`` once."""
    loss, metrics, cfg, coefs, input_size, has_value_head = _forward_loss(model, batch, current_step, max_step, config,
                                                                          vocab_group, v_offset)
    m = metrics.tolist()  # the one host sync of the step
    stats = stats_from_metrics(m, cfg, input_size, has_value_head, coefs)
    return loss, stats


def batch_sequence_count(batch) -> int:
    """finetune_loop.py:266-276 (``get_batch_sequence_count``)."""
    get = (lambda k: batch.get(k)) if isinstance(batch, dict) else (lambda k: getattr(batch, k, None))
    if get("position_ids") is not None and get("seq_boundaries") is not None:
        return len(get("seq_boundaries")) - (1 if not get("padding") else 2)
    return int(get("input_ids").shape[0])


def rl_step_async(model, batch, current_step: int, max_step: int, config, step_metrics, record: bool = True, *,
                  vocab_group=None, v_offset: int | None = None) -> torch.Tensor:
    """``rl_step`` without its host sync: the stats go into ``step_metrics`` (metrics.StepMetrics) on the device.

    Replaces ``loss, m = rl_step(...); for k, v in m.items(): rl_metrics[k].append(v)`` of
    finetune_loop.py:661-670; ``record=False`` is the sentinel-batch branch (:664-666) -- the caller still
    multiplies the loss by 0.  The reference's asserts fire at ``step_metrics.reduce``.
    """
    loss, metrics, cfg, coefs, input_size, _ = _forward_loss(model, batch, current_step, max_step, config, vocab_group, v_offset)
    step_metrics.add(metrics, coefs[0], coefs[1], input_size, batch_sequence_count(batch), record)
    return loss
