"""RLConfig with the reference's field names, defaults and validation behaviour.

Mirror of pipelinerl/finetune/rl/__init__.py:38-99: a pydantic model, so unknown YAML keys
(``aggregate_loss`` ...) are ignored exactly as in the reference, and ``reward_minus_kl_coef``
is accepted and never read.  ``rl_step`` only uses attribute access, so the reference's own
RLConfig instance works as well (duck typing).
"""
from __future__ import annotations

from pydantic import BaseModel, Field


class RLConfig(BaseModel):
    policy_loss: str = Field(default="ppo", description="Policy Loss to use for RL")
    use_advantages: bool = Field(default=True)
    epsilon: float = Field(default=0.2)
    batch_size: int = Field(default=0)
    reward_minus_kl_coef: float = Field(default=0.0)
    kl_coef: float = Field(default=0.1)
    final_kl_coef: float = Field(default=0.1)
    entropy_bonus: float = Field(default=0.0)
    final_entropy_bonus: float = Field(default=0.0)
    relu_log_p_weights: bool = Field(default=False)
    clamp_log_ratio_ref_new_value: float = Field(default=10)
    divide_advantage_by_std: bool = Field(default=True)
    overlong_filtering: bool = Field(default=False)
    group_normalization: bool = Field(default=False)
    temperature: float = Field(default=1.0)
    filter_zero_advantage_groups: bool = Field(default=False)
    value_loss_coef: float = Field(default=0.0)


RL_DATA_COLUMNS = [  # pipelinerl/finetune/rl/__init__.py:27-35 (order matters to collate)
    "overflow",
    "group_tokens",
    "num_labels",
    "rewards",
    "advantages",
    "old_logprobs",
    "ref_logprobs",
]


def linear_decay_coef(current_step: int, max_step: int, initial_coef: float, final_coef: float) -> float:
    """pipelinerl/finetune/rl/__init__.py:113-127 (host scalar, fp64 like the reference)."""
    return initial_coef + (final_coef - initial_coef) * current_step / max_step
