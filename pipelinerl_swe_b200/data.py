"""Packed-batch layout producers: collate_packed / create_sentinel_batch (host side).

Same outputs as pipelinerl/finetune/data.py:215-279 and finetune/utils.py:16-75; built
with numpy concatenation instead of one torch.tensor() per example.
"""
from __future__ import annotations

from typing import Any

import numpy as np
import torch

from .config import RL_DATA_COLUMNS
from .types import PipelineBatchEncoding

MASKED_TOKEN_ID = -100


def create_sentinel_example(n_tokens: int, tokenizer=None, model_version: int = 0) -> dict:
    """finetune/utils.py:59-75."""
    eos = tokenizer.eos_token_id
    return {
        "input_ids": n_tokens * [eos], "attention_mask": n_tokens * [1], "labels": n_tokens * [MASKED_TOKEN_ID],
        "position_ids": list(range(n_tokens)),
        "rewards": n_tokens * [0.0], "advantages": n_tokens * [0.0], "ref_logprobs": n_tokens * [0.0],
        "old_logprobs": n_tokens * [0.0], "group_tokens": n_tokens * [1.0], "num_labels": n_tokens * [1.0],
        "overflow": n_tokens * [0.0], "model_version": model_version,
    }


def create_sentinel_batch(device, tokenizer=None, model_version: int = 0) -> PipelineBatchEncoding:
    """finetune/utils.py:16-56: 8 EOS tokens, all labels masked, sentinel=True, packed."""
    eos = getattr(tokenizer, "eos_token_id", 2) if tokenizer else 2
    n = 8
    f = lambda x: torch.full((1, n), x, dtype=torch.float)  # noqa: E731
    batch = PipelineBatchEncoding(
        input_ids=torch.full((1, n), eos, dtype=torch.long), attention_mask=torch.ones(1, n, dtype=torch.long),
        labels=torch.full((1, n), MASKED_TOKEN_ID, dtype=torch.long), position_ids=torch.arange(n)[None],
        rewards=f(0.0), advantages=f(0.0), ref_logprobs=f(0.0), old_logprobs=f(0.0), group_tokens=f(1.0),
        num_labels=f(1.0), overflow=f(0.0), seq_boundaries=torch.tensor([0, n], dtype=torch.int),
        model_version=model_version, sentinel=True, is_packed=True)
    return batch.to_device(device) if device is not None else batch


def collate_packed(examples: list[dict[str, Any]], tokenizer, seq_parallel: int,
                   label_pad_value: int = MASKED_TOKEN_ID) -> PipelineBatchEncoding:
    """finetune/data.py:215-279: one packed [1,T] row; T padded to a multiple of seq_parallel."""
    total = sum(len(e["input_ids"]) for e in examples)
    padding = 0
    if total % seq_parallel != 0:
        padding = seq_parallel - (total % seq_parallel)
        mv = max(e["model_version"] for e in examples)
        examples = examples + [create_sentinel_example(padding, tokenizer=tokenizer, model_version=mv)]
    lens = np.asarray([len(e["input_ids"]) for e in examples], dtype=np.int64)
    bounds = np.zeros(len(examples) + 1, dtype=np.int64)
    np.cumsum(lens, out=bounds[1:])
    T = int(bounds[-1])
    ids = np.concatenate([np.asarray(e["input_ids"], dtype=np.int64) for e in examples]) if T else np.zeros(0, np.int64)
    labels = np.concatenate([np.asarray(e["labels"], dtype=np.int64) for e in examples]) if T else np.zeros(0, np.int64)
    first = bounds[1:-1][lens[1:] > 0]  # first token of every example after the first (:260-261)
    labels[first] = label_pad_value
    pos = np.arange(T, dtype=np.int64) - np.repeat(bounds[:-1], lens)
    result: dict[str, Any] = {
        "input_ids": torch.from_numpy(ids)[None], "labels": torch.from_numpy(labels)[None],
        "attention_mask": torch.ones(1, T, dtype=torch.long), "position_ids": torch.from_numpy(pos)[None],
    }
    for col in RL_DATA_COLUMNS:
        if col in examples[0]:
            parts = [np.asarray(e[col], dtype=np.float64).reshape(-1) for e in examples]
            result[col] = torch.from_numpy(np.concatenate(parts).astype(np.float32))[None]
    result["model_version"] = min(e.get("model_version", 0) for e in examples)
    result["is_packed"] = True
    result["seq_boundaries"] = torch.from_numpy(bounds.astype(np.int32))
    result["padding"] = padding
    return PipelineBatchEncoding(**result)
