"""pipelinerl_swe_b200: the trainer-side GRPO hot path of ServiceNow/PipelineRL-SWE on B200.

Drop-in names (same signatures as pipelinerl.finetune.rl / finetune.types / finetune.data):
``rl_step``, ``populate_rl_data``, ``prepare_rl_fields``, ``RLConfig``, ``RL_DATA_COLUMNS``,
``aggregate_rl_stats``, ``PipelineBatchEncoding``, ``collate_packed``,
``create_sentinel_batch``.  Compute is in libb200grpo.so (hand-written sm_100a CUDA behind
the C ABI of include/b200grpo.h); importing this package does not need a GPU, calling it does.
"""
from .config import RL_DATA_COLUMNS, RLConfig, linear_decay_coef  # noqa: F401
from .types import PipelineBatchEncoding  # noqa: F401
from .stats import aggregate_rl_stats  # noqa: F401
from .data import collate_packed, create_sentinel_batch, create_sentinel_example  # noqa: F401
from .advantage import filter_zero_advantage_groups, populate_rl_data, prepare_rl_fields  # noqa: F401
from .lm_head import (FusedLMHead, fold_main_grad, install_fused_lm_head, lmhead_logprobs, set_fused_grad_accumulation,  # noqa: F401
                      set_save_logits, set_saved_logits_budget, set_skip_masked_tokens)
from .final_norm import install_fused_final_norm  # noqa: F401
from .rl import rl_step, rl_step_async  # noqa: F401
from .metrics import StepMetrics  # noqa: F401
from .ingest import PackedWire, ingest_packed, pack_rollouts  # noqa: F401
from .ref_logprobs import annotate_traces_with_ref_logprobs  # noqa: F401
from . import weight_update  # noqa: F401
from . import host_step  # noqa: F401

__all__ = [
    "RLConfig", "RL_DATA_COLUMNS", "linear_decay_coef", "PipelineBatchEncoding", "aggregate_rl_stats", "collate_packed",
    "create_sentinel_batch", "create_sentinel_example", "populate_rl_data", "prepare_rl_fields",
    "filter_zero_advantage_groups", "FusedLMHead", "install_fused_lm_head", "install_fused_final_norm", "lmhead_logprobs", "set_skip_masked_tokens",
    "set_save_logits", "set_saved_logits_budget", "set_fused_grad_accumulation", "fold_main_grad", "rl_step", "rl_step_async", "StepMetrics", "PackedWire", "ingest_packed", "pack_rollouts",
    "annotate_traces_with_ref_logprobs", "weight_update", "host_step",
]
