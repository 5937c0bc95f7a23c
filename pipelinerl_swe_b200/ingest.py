"""Packed-batch ingest: 16 B/token + 20 B/rollout on the wire, expanded on the GPU.

Reference behaviour replaced: ``collate_packed`` (pipelinerl/finetune/data.py:215-279) materialises eleven
[1, T] host tensors -- int64 ids / labels / attention_mask / position_ids and seven float columns, 60 B per
token -- which ``PipelineBatchEncoding.to_device`` (finetune_loop.py:109, types.py:111-117) copies one by
one; five of the float columns (rewards, advantages, group_tokens, num_labels, overflow) are per-ROLLOUT
values that ``populate_rl_data`` broadcast to every token on the CPU (rl/__init__.py:481-499).

``pack_rollouts`` writes the wire form into ONE pinned buffer (struct of arrays):

    input_ids i32[T] | labels i32[T] | old_logprobs f32[T] | ref_logprobs f32[T]
    | seq_boundaries i32[n+1] | rewards, advantages, group_tokens, num_labels, overflow f32[5][n]

``PackedWire.to_device`` issues one H2D copy and one kernel (``b200grpo_expand_packed``) that produces the
same eleven tensors ``collate_packed(...).to_device(...)`` would hold -- bit for bit -- so ``rl_step``
needs no change.  Examples whose per-rollout columns are not constant inside a rollout (per-token
rewards) cannot take the wire form; ``ingest_packed`` falls back to ``collate_packed`` for them.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass
from typing import Any

import numpy as np
import torch

from . import _abi
from .config import RL_DATA_COLUMNS
from .data import MASKED_TOKEN_ID, collate_packed, create_sentinel_example
from .types import PipelineBatchEncoding

SEQ_COLUMNS = ("rewards", "advantages", "group_tokens", "num_labels", "overflow")  # row order of seq_scalars
_ALIGN = 16


class NonUniformColumns(ValueError):
    """A per-rollout column varies inside a rollout: the batch cannot take the 20 B/rollout wire form."""


def _layout(T: int, n: int) -> tuple[dict[str, tuple[int, int]], int]:
    """name -> (byte offset, byte length); every section 16-byte aligned."""
    sections = [("input_ids", 4 * T), ("labels", 4 * T), ("old_logprobs", 4 * T), ("ref_logprobs", 4 * T),
                ("seq_boundaries", 4 * (n + 1)), ("seq_scalars", 4 * 5 * n)]
    out, off = {}, 0
    for name, nbytes in sections:
        out[name] = (off, nbytes)
        off += (nbytes + _ALIGN - 1) // _ALIGN * _ALIGN
    return out, max(off, _ALIGN)


@dataclass
class PackedWire:
    buffer: torch.Tensor          # uint8, host (pinned when a GPU is present)
    T: int
    n_seq: int
    model_version: int
    padding: int = 0
    sentinel: bool = False

    @property
    def nbytes(self) -> int:
        return int(self.buffer.numel())

    def section(self, name: str, buf: torch.Tensor | None = None) -> torch.Tensor:
        lay, _ = _layout(self.T, self.n_seq)
        off, n = lay[name]
        dt = torch.float32 if name in ("old_logprobs", "ref_logprobs", "seq_scalars") else torch.int32
        return (self.buffer if buf is None else buf)[off:off + n].view(dt)

    def to_device(self, device: torch.device | str) -> PipelineBatchEncoding:
        """One H2D copy + one kernel -> the batch ``collate_packed(...).to_device(device)`` would give."""
        dev = torch.device(device)
        if dev.type != "cuda":
            raise _abi.B200GrpoError("PackedWire.to_device: CUDA device required (no CPU path)")
        T, n = self.T, self.n_seq
        dbuf = torch.empty(self.nbytes, dtype=torch.uint8, device=dev)
        dbuf.copy_(self.buffer, non_blocking=True)
        i64 = torch.empty(4, max(T, 1), dtype=torch.int64, device=dev)
        f32 = torch.empty(7, max(T, 1), dtype=torch.float32, device=dev)
        cols = ("rewards", "advantages", "ref_logprobs", "old_logprobs", "group_tokens", "num_labels", "overflow")
        out = _abi.PackedOut(*[i64[k].data_ptr() for k in range(4)], *[f32[k].data_ptr() for k in range(7)])
        sec = lambda name: self.section(name, dbuf).data_ptr()  # noqa: E731
        _abi.check(_abi.lib().b200grpo_expand_packed(
            sec("input_ids"), sec("labels"), sec("old_logprobs"), sec("ref_logprobs"), sec("seq_boundaries"), sec("seq_scalars"),
            T, n, C.byref(out), torch.cuda.current_stream(dev).cuda_stream))
        fields: dict[str, Any] = {k: i64[j, :T][None] for j, k in enumerate(("input_ids", "labels", "attention_mask", "position_ids"))}
        fields.update({k: f32[j, :T][None] for j, k in enumerate(cols)})
        return PipelineBatchEncoding(**fields, model_version=self.model_version, sentinel=self.sentinel, padding=self.padding,
                                     is_packed=True, seq_boundaries=self.section("seq_boundaries", dbuf))


def _seq_scalar(example: dict[str, Any], col: str, n_tokens: int) -> float:
    v = example[col]
    if isinstance(v, np.ndarray):
        v = v.reshape(-1).tolist()
    if isinstance(v, (list, tuple)):
        n = len(v)
        if n == 0:
            return 0.0
        if n != n_tokens:
            raise NonUniformColumns(f"column {col}: {n} values for {n_tokens} tokens")
        if v.count(v[0]) != n:       # C-speed scan, no array conversion
            raise NonUniformColumns(f"column {col} varies inside a rollout")
        return float(v[0])
    return float(v)


def pack_rollouts(examples: list[dict[str, Any]], tokenizer=None, seq_parallel: int = 1,
                  label_pad_value: int = MASKED_TOKEN_ID, pin: bool | None = None) -> PackedWire:
    """Wire form of ``collate_packed(examples, tokenizer, seq_parallel)`` (same padding / label rules)."""
    missing = [c for c in RL_DATA_COLUMNS if c not in examples[0]]
    if missing:
        raise NonUniformColumns(f"examples lack RL columns {missing}")
    total = sum(len(e["input_ids"]) for e in examples)
    padding = 0
    if total % seq_parallel != 0:  # data.py:222-228
        padding = seq_parallel - (total % seq_parallel)
        mv = max(e["model_version"] for e in examples)
        examples = examples + [create_sentinel_example(padding, tokenizer=tokenizer, model_version=mv)]
    n = len(examples)
    lens = np.asarray([len(e["input_ids"]) for e in examples], dtype=np.int64)
    bounds = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(lens, out=bounds[1:])
    T = int(bounds[-1])
    if T >= 2 ** 31:
        raise ValueError("packed row too long")
    _, nbytes = _layout(T, n)
    pin = torch.cuda.is_available() if pin is None else pin
    buf = torch.empty(nbytes, dtype=torch.uint8, pin_memory=bool(pin))
    wire = PackedWire(buf, T, n, min(e.get("model_version", 0) for e in examples), padding)
    ids, labels = wire.section("input_ids").numpy(), wire.section("labels").numpy()
    old, ref = wire.section("old_logprobs").numpy(), wire.section("ref_logprobs").numpy()
    scal = wire.section("seq_scalars").numpy().reshape(5, n)
    for i, e in enumerate(examples):
        a, b = int(bounds[i]), int(bounds[i + 1])
        ids[a:b] = e["input_ids"]
        labels[a:b] = e["labels"]
        if i > 0 and b > a:
            labels[a] = label_pad_value  # data.py:260-261
        old[a:b] = np.asarray(e["old_logprobs"], dtype=np.float64)
        ref[a:b] = np.asarray(e["ref_logprobs"], dtype=np.float64)
        for r, col in enumerate(SEQ_COLUMNS):
            scal[r, i] = _seq_scalar(e, col, b - a)
    wire.section("seq_boundaries").numpy()[:] = bounds
    return wire


def ingest_packed(examples: list[dict[str, Any]], tokenizer, seq_parallel: int, device: torch.device | str,
                  label_pad_value: int = MASKED_TOKEN_ID) -> PipelineBatchEncoding:
    """``collate_packed(...).to_device(device)`` through the wire form when the examples allow it."""
    try:
        wire = pack_rollouts(examples, tokenizer, seq_parallel, label_pad_value)
    except NonUniformColumns:
        return collate_packed(examples, tokenizer, seq_parallel, label_pad_value).to_device(device)
    return wire.to_device(device)
