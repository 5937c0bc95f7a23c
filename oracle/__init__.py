"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference's GRPO hot path.

Nothing under ``oracle/`` is part of the product.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference``
legs may import it, and only as the checker / the timed CPU baseline.  The product
package (``pipelinerl_swe_b200``) never imports this module and fails loudly when its
CUDA library is missing.

Parity pinning: the reference (ServiceNow/PipelineRL-SWE) ships no tests, fixtures or
golden vectors for this path.  The oracle is therefore
pinned against outputs of the *reference itself*, produced in the authoring container
by ``tests/golden/make_golden.py`` (which imports ``/root/reference`` unmodified) and
committed as ``tests/golden/*.npz``; ``tests/test_oracle_golden.py`` checks every
oracle function against them.
"""
from .grpo_oracle import (  # noqa: F401
    OracleRLConfig,
    oracle_aggregate_rl_stats,
    oracle_collate_packed,
    oracle_filter_zero_advantage_groups,
    oracle_final_norm_value,
    oracle_linear_decay_coef,
    oracle_logprobs_entropy,
    oracle_merge_stats,
    oracle_populate_rl_data,
    oracle_prepare_rl_fields,
    oracle_rl_step,
    oracle_segments,
    oracle_sentinel_batch,
    oracle_shard_stats,
    oracle_token_loss,
)
