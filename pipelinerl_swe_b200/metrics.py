"""Fixed-layout step metrics: device accumulation + ONE collective per optimizer step.

Reference behaviour replaced (pipelinerl/finetune_loop.py):
  * :669-670  ``rl_metrics[k].append(v)`` for every key of every micro-batch (python floats, which is
              why ``rl_step`` must sync the GPU once per micro-batch);
  * :62-87    ``gather_rl_metrics``: ``dist.all_gather_object`` of the pickled dict of lists, non-finite
              values dropped;
  * :767      ``aggregate_rl_stats`` (rl/utils.py:8-22): reduction picked by SUBSTRING of the key
              ("min" -> min, "max" -> max, "loss" == key or "sum" -> sum, else sum / num_samples).

Here each micro-batch folds its K3 metrics vector into a device vector with one tiny kernel and no host
sync (``StepMetrics.add``); at the optimizer-step boundary ``StepMetrics.reduce`` moves
``2K + 4`` doubles per rank in one ``all_gather_into_tensor`` and combines them on the host.  The
reference's asserts on non-finite log-probs / loss travel as a flag word and fire at ``reduce``.
"""
from __future__ import annotations

from typing import Any

import torch

from . import _abi

A = _abi

# key -> metric slot, in the order of the reference's stats dict (rl/__init__.py:325-375)
STAT_SLOTS: list[tuple[str, int]] = [
    ("loss", A.M_LOSS), ("max_loss", A.M_LOSS), ("min_loss", A.M_LOSS),
    ("reward", A.M_REWARD), ("max_reward", A.M_MAX_REWARD), ("min_reward", A.M_MIN_REWARD),
    ("entropy", A.M_ENTROPY), ("old_logprobs", A.M_OLD_LP), ("new_logprobs", A.M_NEW_LP), ("ref_logprobs", A.M_REF_LP),
    ("advantage", A.M_ADVANTAGE), ("max_advantage", A.M_MAX_ADV), ("min_advantage", A.M_MIN_ADV),
    ("kl", A.M_KL), ("max_kl", A.M_MAX_KL), ("min_kl", A.M_MIN_KL),
    ("policy_loss", A.M_POLICY), ("surr1", A.M_SURR1), ("surr2", A.M_SURR2),
    ("ratio_new_old", A.M_RATIO), ("ratio_new_old_sum", A.M_RATIO_SUM), ("ratio_new_old_squared_sum", A.M_RATIO_SQ_SUM),
    ("ratio_ref_new", A.M_RATIO_REF_NEW), ("ratio_ref_old", A.M_RATIO_REF_OLD),
    ("clamp_log_ratio_ref_new_indicator", A.M_CLAMP_REF_NEW), ("clamp_log_ratio_new_old_indicator", A.M_CLAMP_NEW_OLD),
    ("num_nans", A.M_NUM_NANS),
    ("token_weight", A.M_TOKEN_WEIGHT), ("max_token_weight", A.M_MAX_TW), ("min_token_weight", A.M_MIN_TW),
    ("kl_coef", A.STAT_KL_COEF), ("entropy_bonus_coef", A.STAT_ENTROPY_COEF),
    ("num_output_tokens_sum", A.M_N_MASKED), ("input_size", A.STAT_INPUT_SIZE),
]
VALUE_STAT_SLOTS: list[tuple[str, int]] = [
    ("value_mean", A.M_VALUE_MEAN), ("value_max", A.M_MAX_VALUE), ("value_min", A.M_MIN_VALUE),
    ("value_loss", A.M_VALUE_LOSS), ("value_mse", A.M_VALUE_MSE),
]
KIND_SUM, KIND_MIN, KIND_MAX, KIND_MEAN = 0, 1, 2, 3


def stat_kind(key: str) -> int:
    """rl/utils.py:10-20, same test order."""
    if "min" in key:
        return KIND_MIN
    if "max" in key:
        return KIND_MAX
    if key == "loss" or "sum" in key:
        return KIND_SUM
    return KIND_MEAN


def stat_layout(has_values: bool) -> tuple[list[str], list[int], list[int]]:
    """(keys, slots, kinds) of the accumulator."""
    table = STAT_SLOTS + (VALUE_STAT_SLOTS if has_values else [])
    keys = [k for k, _ in table]
    return keys, [s for _, s in table], [stat_kind(k) for k in keys]


def combine(acc: torch.Tensor, kinds: list[int]) -> torch.Tensor:
    """[R, 2K+4] per-rank accumulators -> [2K+4]: values by kind, counts / counters summed, flags or-ed."""
    K = len(kinds)
    acc = acc.to(torch.float64).reshape(-1, 2 * K + 4)
    out = acc.sum(0)
    kind_t = torch.tensor(kinds)
    vals = acc[:, :K]
    out[:K] = torch.where(kind_t == KIND_MIN, vals.min(0).values, torch.where(kind_t == KIND_MAX, vals.max(0).values, vals.sum(0)))
    flags = 0
    for f in acc[:, 2 * K].tolist():
        flags |= int(f)
    out[2 * K] = float(flags)
    return out


def check_flags(flags: int) -> None:
    """The reference's in-step asserts (rl/__init__.py:209,222,237,264,313), raised at the step boundary."""
    assert not flags & A.F_GROUP_TOKENS_LE0, "group_tokens must be greater than zero for group normalization"
    assert not flags & A.F_NEW_LP_NONFINITE, "new_logprobs is not finite"
    assert not flags & A.F_REF_NEW_NONFINITE, "log_ratio_ref_new is not finite"
    assert not flags & A.F_KL_NONFINITE, "approx_kl is not finite"
    assert not flags & A.F_LOSS_NONFINITE, "Non-finite loss detected"


def finalize(total: torch.Tensor, keys: list[str], kinds: list[int], num_samples: int) -> dict[str, float]:
    """Combined accumulator -> the dict ``aggregate_rl_stats(gather_rl_metrics(...), num_samples)`` returns.

    Keys that never received a finite value are absent, like keys with empty lists in the reference (:84)."""
    K = len(keys)
    t = total.tolist()
    out: dict[str, float] = {}
    for i, k in enumerate(keys):
        if t[K + i] == 0:
            continue
        out["rl/" + k] = t[i] / num_samples if kinds[i] == KIND_MEAN else t[i]
    return out


def effective_sample_size(avg: dict[str, float]) -> float:
    """finetune_loop.py:768-772."""
    return avg["rl/ratio_new_old_sum"] ** 2 / avg["rl/ratio_new_old_squared_sum"] / avg["rl/num_output_tokens_sum"]


class StepMetrics:
    """Device-side accumulator of one optimizer step's RL metrics.

        sm = StepMetrics(device, has_values=False)
        for micro_batch in ...:
            loss = rl_step_async(model, batch, step, max_step, cfg, sm)     # no host sync
            loss.backward()
        avg, counters = sm.reduce(samples_per_step, group)                  # one collective + one D2H
    """

    def __init__(self, device: torch.device | str, has_values: bool = False):
        self.device = torch.device(device)
        self.keys, slots, self.kinds = stat_layout(has_values)
        self.K = len(self.keys)
        assert self.K <= A.STATS_MAX
        self._slots = torch.tensor(slots, dtype=torch.int32, device=self.device)
        abi_kinds = [A.STAT_MIN if k == KIND_MIN else A.STAT_MAX if k == KIND_MAX else A.STAT_SUM for k in self.kinds]
        self._kinds = torch.tensor(abi_kinds, dtype=torch.int32, device=self.device)
        self.acc = torch.empty(2 * self.K + 4, dtype=torch.float64, device=self.device)
        self.reset()

    def reset(self) -> None:
        _abi.check(_abi.lib().b200grpo_metrics_reset(self._kinds.data_ptr(), self.K, self.acc.data_ptr(), _stream(self.device)))

    def add(self, metrics: torch.Tensor, kl_coef: float, entropy_coef: float, input_size: int, n_samples: int = 0,
            record: bool = True) -> None:
        """Fold one micro-batch (K3's f64[40] metrics, still on the device).  ``record=False``: sentinel batch."""
        assert metrics.dtype == torch.float64 and metrics.numel() >= _abi.M_COUNT and metrics.device == self.device
        _abi.check(_abi.lib().b200grpo_metrics_accumulate(
            metrics.data_ptr(), self._slots.data_ptr(), self._kinds.data_ptr(), self.K, float(kl_coef), float(entropy_coef),
            float(input_size), float(n_samples), int(record), self.acc.data_ptr(), _stream(self.device)))

    def reduce(self, num_samples: int, group: Any = None, reset: bool = True) -> tuple[dict[str, float], dict[str, int]]:
        """-> (``rl/...`` averages as the reference logs them, counters).  One collective when ``group`` has > 1 rank."""
        avg, counters = reduce_accumulators(self.acc, self.keys, self.kinds, num_samples, group)
        if reset:
            self.reset()
        return avg, counters


def reduce_accumulators(acc: torch.Tensor, keys: list[str], kinds: list[int], num_samples: int, group: Any = None):
    """Exchange + combine + finalize; device-agnostic (the exchange is what the gloo tests cover)."""
    import torch.distributed as dist
    world = dist.get_world_size(group) if (group is not None or (dist.is_available() and dist.is_initialized())) else 1
    if world > 1:
        gathered = torch.empty(world * acc.numel(), dtype=acc.dtype, device=acc.device)
        dist.all_gather_into_tensor(gathered, acc.contiguous(), group=group)
    else:
        gathered = acc
    K = len(keys)
    total = combine(gathered.cpu(), kinds)  # the step's one D2H copy
    check_flags(int(total[2 * K]))
    counters = {"labelled_micro_batches": int(total[2 * K + 1]), "micro_batches": int(total[2 * K + 2]), "samples": int(total[2 * K + 3])}
    return finalize(total, keys, kinds, num_samples), counters


def _stream(device: torch.device) -> int:
    return torch.cuda.current_stream(device).cuda_stream
