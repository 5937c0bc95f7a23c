"""Advantage step: populate_rl_data / prepare_rl_fields / filter_zero_advantage_groups.

Drop-ins for pipelinerl/finetune/rl/__init__.py:380-525 and preprocess.py:287-324 with the
same list-of-dict in/out contract.  The pandas groupby/merge/apply of the reference is
replaced by: per-ROLLOUT scalars gathered on the host (reward, dense group key, length,
label count, has-eos; no per-token Python work), one K4 launch on the GPU (group statistics in
f64 + one value per rollout and column, or the four per-token columns when rewards vary inside
a rollout), one D2H copy.
"""
from __future__ import annotations

from typing import Any

import numpy as np
import torch

from . import ops

IGNORE = -100


def prepare_rl_fields(encoding: dict[str, Any], reward: float, old_logprobs: list[float], ref_logprobs: list[float]) -> dict[str, Any]:
    """rl/__init__.py:504-525 (pure host list bookkeeping; nothing to accelerate)."""
    n = len(encoding["labels"])
    labels = encoding["labels"]
    n_target = (len(labels) - labels.count(IGNORE)) if isinstance(labels, list) else sum(1 for t in labels if t != IGNORE)
    assert n_target == len(old_logprobs), f"Target tokens: {n_target}, old logprobs: {len(old_logprobs)}"
    encoding["rewards"] = [reward] * n
    encoding["advantages"] = [0.0] * n
    encoding["old_logprobs"] = [0] * (n - len(old_logprobs)) + old_logprobs
    encoding["ref_logprobs"] = [0] * (n - len(ref_logprobs)) + ref_logprobs
    encoding["overflow"] = [0] * n
    encoding["group_tokens"] = [0] * n
    encoding["num_labels"] = [1 if t != IGNORE else 0 for t in encoding["labels"]]
    return encoding


def rollout_scalars(dataset: list[dict[str, Any]], eos_token_id: int):
    """Host side of K4's input: one record per rollout row (17 B/rollout).  Per-row scans are C-speed list methods
    (``count`` / ``in``), not Python loops or array conversions: 4096 rollouts x 2 k tokens take ~30 ms."""
    keys: dict[str, int] = {}
    n = len(dataset)
    reward = np.empty(n, dtype=np.float64)
    gkey = np.empty(n, dtype=np.int32)
    lens = np.empty(n, dtype=np.int64)
    nlab = np.empty(n, dtype=np.int32)
    eos = np.empty(n, dtype=np.uint8)
    seen: dict[tuple[int, Any], float] = {}
    uniform = True
    for i, row in enumerate(dataset):
        k = f"{row['group_id']}_{row['step_index']}"  # :395
        gkey[i] = keys.setdefault(k, len(keys))
        r = row["rewards"]
        r0 = r[0]
        reward[i] = r0  # :402
        # :405 -- one reward per (group, rollout)
        assert seen.setdefault((int(gkey[i]), row["rollout_index"]), r0) == r0
        if uniform:
            r_list = r if isinstance(r, list) else list(r)
            uniform = r_list.count(r0) == len(r_list)
        ids, labels = row["input_ids"], row["labels"]
        ids = ids if isinstance(ids, list) else list(ids)
        labels = labels if isinstance(labels, list) else list(labels)
        lens[i] = len(ids)
        nlab[i] = len(labels) - labels.count(IGNORE)
        eos[i] = eos_token_id in ids
    offsets = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(lens, out=offsets[1:])
    # stable sort of the rows by group key (numpy radix argsort over int32: O(n)) -> the CSR the stats kernel walks
    order = np.argsort(gkey, kind="stable").astype(np.int32)
    starts = np.zeros(len(keys) + 1, dtype=np.int32)
    np.cumsum(np.bincount(gkey, minlength=len(keys)), out=starts[1:])
    tok_rewards = None
    if not uniform:  # per-token rewards differ inside a row: ship them (f64) so :466 is followed literally
        tok_rewards = np.concatenate([np.asarray(row["rewards"], dtype=np.float64) for row in dataset])
    return reward, gkey, offsets, nlab, eos, len(keys), tok_rewards, order, starts


def populate_rl_data(dataset: list[dict[str, Any]], eos_token_id: int, config, device: torch.device | str | None = None) -> list[dict[str, Any]]:
    """rl/__init__.py:380-501: fills advantages / group_tokens / overflow / num_labels per token, in place.

    Rewards constant inside every rollout (the reference's own producers, preprocess.py:163 via prepare_rl_fields):
    K4 returns one value per rollout and column (16 B/rollout back), the per-token lists are list multiplications.
    Per-token rewards: K4 expands on the GPU (16 B/token back) so that :466 is followed literally."""
    if not dataset:
        return dataset
    dev = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
    reward, gkey, offsets, nlab, eos, n_groups, tok_rewards, order, starts = rollout_scalars(dataset, eos_token_id)
    to = lambda a: torch.from_numpy(a).to(dev, non_blocking=True)  # noqa: E731
    std = bool(config.divide_advantage_by_std)
    if tok_rewards is None:
        rows = ops.group_advantage_rows(to(reward), to(gkey), to(offsets), n_groups, std, to(order), to(starts))
        adv, gtok = rows.cpu().numpy().tolist()  # f64 -> Python floats: the very values the reference's lists hold (:461-469, :486)
        for i, row in enumerate(dataset):
            n = int(offsets[i + 1] - offsets[i])
            row["advantages"] = [adv[i]] * len(row["rewards"])
            row["group_tokens"] = [gtok[i]] * n
            row["overflow"] = [0.0 if eos[i] else 1.0] * len(row["overflow"])   # :481-484
            row["num_labels"] = [int(nlab[i])] * n
        return dataset
    # rewards vary inside a rollout: the GPU forms every token's f64 advantage (:466 followed literally)
    out, adv64 = ops.group_advantage(to(reward), to(gkey), to(offsets), to(nlab), to(eos), n_groups, std, to(tok_rewards),
                                     to(order), to(starts), want_f64=True)
    host = adv64.cpu().numpy()
    gtok = ops.group_advantage_rows(to(reward), to(gkey), to(offsets), n_groups, std, to(order), to(starts))[1].cpu().numpy().tolist()
    for i, row in enumerate(dataset):
        a, b = int(offsets[i]), int(offsets[i + 1])
        n_adv = len(row["rewards"])
        row["advantages"] = host[a:a + n_adv].tolist()
        row["group_tokens"] = [gtok[i]] * (b - a)
        row["overflow"] = [0.0 if eos[i] else 1.0] * len(row["overflow"])
        row["num_labels"] = [int(nlab[i])] * (b - a)
    return dataset


def filter_zero_advantage_groups(dataset: list[dict], epsilon: float = 1e-6) -> tuple[list[dict], int]:
    """preprocess.py:287-324: drop groups whose advantages are all within epsilon of zero."""
    groups: dict[Any, list[dict]] = {}
    for entry in dataset:
        groups.setdefault(entry["group_id"], []).append(entry)
    kept: list[dict] = []
    dropped = 0
    for entries in groups.values():
        if any(bool((np.abs(np.asarray(e["advantages"], dtype=np.float64)) > epsilon).any()) for e in entries):
            kept.extend(entries)
        else:
            dropped += len(entries)
    return kept, dropped
