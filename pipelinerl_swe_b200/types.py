"""PipelineBatchEncoding: the batch contract of pipelinerl/finetune/types.py:48-178.

Same field names, dtype coercions (long / int32 / float32), ``to_device`` and
``make_slices`` semantics, so stream payloads written by the reference's preprocessor
deserialise into this class unchanged (``PipelineBatchEncoding(**dict)``).
"""
from __future__ import annotations

from typing import Any, Dict, List, Union

import numpy as np
import torch
from pydantic import BaseModel, ConfigDict, field_validator

_SLICED = ("input_ids", "attention_mask", "labels", "position_ids", "rewards", "advantages", "ref_logprobs", "old_logprobs",
           "group_tokens", "overflow", "num_labels")


class PipelineBatchEncoding(BaseModel):
    model_config = ConfigDict(arbitrary_types_allowed=True)

    input_ids: torch.Tensor
    attention_mask: torch.Tensor
    labels: torch.Tensor
    position_ids: torch.Tensor | None = None
    rewards: torch.Tensor
    advantages: torch.Tensor
    ref_logprobs: torch.Tensor
    old_logprobs: torch.Tensor
    group_tokens: torch.Tensor
    num_labels: torch.Tensor
    overflow: torch.Tensor
    model_version: int
    sentinel: bool = False
    padding: int = 0
    is_packed: bool = False
    seq_boundaries: torch.Tensor | None = None
    pixel_values: torch.Tensor | None = None
    image_grid_thw: torch.Tensor | None = None
    # Beyond the reference (opt-in, see make_slices): the first token of the NEXT sequence-parallel slice -- every sliced
    # field as a [1, 1] tensor -- so that this slice's last position can be scored against it.
    boundary: Dict[str, torch.Tensor] | None = None

    @field_validator("input_ids", "attention_mask", "labels", "position_ids", "image_grid_thw", mode="before")
    @classmethod
    def _long(cls, v):
        if v is None:
            return None
        if isinstance(v, torch.Tensor):
            return v.long()
        if isinstance(v, (list, np.ndarray)):
            return torch.tensor(v, dtype=torch.long)
        raise ValueError(f"Unsupported type for long tensor: {type(v)}")

    @field_validator("seq_boundaries", mode="before")
    @classmethod
    def _int(cls, v):
        if v is None:
            return None
        if isinstance(v, torch.Tensor):
            return v.int()
        return torch.tensor(v, dtype=torch.int)

    @field_validator("rewards", "advantages", "ref_logprobs", "old_logprobs", "group_tokens", "num_labels", "overflow",
                     "pixel_values", mode="before")
    @classmethod
    def _float(cls, v):
        if v is None:
            return None
        if isinstance(v, torch.Tensor):
            return v.float()
        if isinstance(v, (list, np.ndarray)):
            return torch.tensor(v, dtype=torch.float)
        raise ValueError(f"Unsupported type for float tensor: {type(v)}")

    def to_device(self, device: Union[str, torch.device]) -> "PipelineBatchEncoding":
        for name in type(self).model_fields:
            value = getattr(self, name)
            if isinstance(value, torch.Tensor):
                setattr(self, name, value.to(device, non_blocking=True))
        if self.boundary is not None:
            self.boundary = {k: v.to(device, non_blocking=True) for k, v in self.boundary.items()}
        return self

    @classmethod
    def from_dict(cls, data: Dict[str, Any], **defaults) -> "PipelineBatchEncoding":
        merged = {**defaults, **data}
        return cls(**{k: v for k, v in merged.items() if k in cls.model_fields})

    def make_slices(self, num_slices: int, carry_boundary_target: bool = False) -> List["PipelineBatchEncoding"]:
        """Contiguous equal cuts of every [1, L] tensor (types.py:144-178).

        The reference scores position t of a slice against token t+1 OF THE SAME SLICE (rl/__init__.py:200,208), so the first
        token of every slice but the first is nobody's target: ``num_slices - 1`` labelled tokens per row drop out of the
        loss.  ``carry_boundary_target=True`` (opt-in, default off = the reference's behaviour bit for bit) attaches to each
        slice but the last the next slice's first token (``boundary``: every sliced field as [1, 1]); ``rl_step`` then scores
        the slice's last position against it -- the model forward is untouched, slices keep equal length (ring attention),
        and the slices together cover exactly the targets of the un-sliced row."""
        if self.position_ids is None or self.input_ids.shape[0] > 1:
            raise ValueError("Cannot a batch that is not properly packed")
        L = self.input_ids.shape[1]
        if L < num_slices:
            raise ValueError(f"Cannot slice batch of size {L} into {num_slices} slices")
        if L % num_slices != 0:
            raise ValueError(f"Sequence length {L} is not divisible by number of slices {num_slices}")
        cuts = [i * L // num_slices for i in range(num_slices + 1)]
        out = []
        for a, b in zip(cuts[:-1], cuts[1:]):
            fields = {k: getattr(self, k)[:, a:b] for k in _SLICED}
            fields.update(model_version=self.model_version, sentinel=self.sentinel, is_packed=self.is_packed,
                          padding=self.padding, seq_boundaries=self.seq_boundaries, pixel_values=self.pixel_values,
                          image_grid_thw=self.image_grid_thw)
            if carry_boundary_target and b < L:
                fields["boundary"] = {k: getattr(self, k)[:, b:b + 1] for k in _SLICED}
            out.append(PipelineBatchEncoding(**fields))
        return out
